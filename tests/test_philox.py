"""The candidate generator's numpy restatement against the published known-answer vectors of Philox4x32-10
(Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", Random123 kat_vectors), and the uniform mapping."""
import numpy as np

from philox_ref import philox4x32_10, uniform_rows


def test_philox_known_answers():
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = philox4x32_10(*[[c] for c in ctr], *key)
        assert tuple(int(v[0]) for v in got) == want


def test_uniform_rows_are_a_pure_function_of_the_index():
    a = uniform_rows(42, 0, 1000, 5, [0, 1, 2, 3, 4], [1, 3, 5, 7, 9])
    b = uniform_rows(42, 300, 50, 5, [0, 1, 2, 3, 4], [1, 3, 5, 7, 9])
    assert np.array_equal(a[300:350], b)
    assert np.all(a >= [0, 1, 2, 3, 4]) and np.all(a < [1, 3, 5, 7, 9])
    assert abs(a[:, 0].mean() - 0.5) < 0.05 and len(np.unique(a)) == a.size
