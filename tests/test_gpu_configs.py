"""GPU parity at the BASELINE.json configs' OWN shapes (tests/parity_cases.py; the achieved numbers are tabulated in
profiles/parity_r02.md by tools/parity_table.py).

Tolerance: the north_star's 1e-10 relative for EI, sd and gradients against the float64 oracle.  Where the oracle itself
(LAPACK's rounding at n = 500 ... 5000) is further than that from the exact value of the reference's formulas, the
extended-precision restatement adjudicates: the GPU must then be within 1e-10 of THAT, and within 5e-10 of the oracle.
The mean and the kriging quantile cross zero, so they are held to 1e-10 of the response scale max|y| + sqrt(sd2).
AKG: a difference of O(1) quantities that nearly cancel (values <= 5e-6 at cfg5): 1e-10 of the response scale; its
float64 evaluation (the reference's own included) keeps about 3 digits of such values (profiles/parity_r02.md)."""
import numpy as np
import pytest

import parity_cases as pc

pytestmark = pytest.mark.gpu

TOL = 1e-10


def _check(rows, loose=5e-10):
    for r in rows:
        a, h = r["gpu_vs_oracle"], r["gpu_vs_hiprec"]
        ok = a <= TOL or (h is not None and h <= TOL and a <= loose)
        assert ok, r
        if h is not None:
            assert h <= TOL, r


def test_cfg1_full_grid_ei_and_grad(gpu_lib):
    rows, info = pc.ei_family("cfg1_branin2_n20", N=None, n_hi=32, n_grad_hi=16, full_grid=True)
    assert rows[0]["points"] == 10_000
    _check(rows)
    assert info["argmax_equal"]


def test_cfg2_shape(gpu_lib):
    rows, info = pc.ei_family("cfg2_hartman4_n200", N=2048, n_hi=32, n_grad_hi=8)
    _check(rows)
    assert info["argmax_equal"]


def test_cfg3_qei_and_qei_grad_256_batches(gpu_lib):
    rows, info = pc.qei_family(B=256, q=4)
    assert info["qei_max"] > 0.1
    _check(rows)


def test_cfg3_predict_family(gpu_lib):
    rows, info = pc.ei_family("cfg3_rosenbrock4_n500", N=1024, n_hi=32, n_grad_hi=8)
    _check(rows)
    assert info["argmax_equal"]


def test_cfg4_shape(gpu_lib):
    rows, info = pc.ei_family("cfg4_sphere10_n2000", N=1024, n_hi=16, n_grad_hi=4)
    _check(rows)
    assert info["argmax_equal"]


@pytest.mark.parametrize("noise_kind", ["noise_var", "nugget"])
def test_cfg5_akg_and_quantile(gpu_lib, noise_kind):
    rows, info = pc.akg_family(noise_kind=noise_kind, N=256, n_hi=16)
    for r in rows:
        if r["quantity"] == "AKG":
            # values only, relative to the largest AKG of the sample (with a homogeneous nugget the slopes
            # are rounding noise, the permutation is not reproducible, the value is)
            assert r["gpu_vs_oracle"] <= (1e-6 if noise_kind == "noise_var" else 1e-8), r
        else:
            _check([r])
    assert np.array_equal(np.isnan(info["akg"]), np.isnan(info["akg_ref"]))
    assert info["akg_max"] > 0 and np.all(info["akg"] >= -1e-12)
