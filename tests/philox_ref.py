"""numpy restatement of the library's counter-based candidate generator (include/dk_b200.h, dk_uniform_spec):
Philox4x32-10, counter = (j lo, j hi, c / 2, 0), key = (seed lo, seed hi); u = ((r_a << 21) | (r_b >> 11)) 2^-53 with
(r_a, r_b) = outputs (0, 1) for even c, (2, 3) for odd c; value = fma(u, upper - lower, lower).  Test infrastructure."""
import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint64) for v in (c0, c1, c2, c3))
    k0 = np.uint64(k0); k1 = np.uint64(k1)
    for _ in range(10):
        p0 = np.uint64(M0) * c0; p1 = np.uint64(M1) * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & MASK; n1 = p1 & MASK
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ k1) & MASK; n3 = p0 & MASK
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(W0)) & MASK; k1 = (k1 + np.uint64(W1)) & MASK
    return c0, c1, c2, c3


def uniform_rows(seed, first, count, d, lower=None, upper=None):
    """rows [first, first + count) of the generated N x d candidate matrix"""
    lower = np.zeros(d) if lower is None else np.asarray(lower, float)
    upper = np.ones(d) if upper is None else np.asarray(upper, float)
    j = np.arange(first, first + count, dtype=np.uint64)
    out = np.empty((count, d))
    for c in range(d):
        r = philox4x32_10(j & MASK, j >> np.uint64(32), np.full(count, c >> 1, np.uint64), np.zeros(count, np.uint64),
                          seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        hi, lo = (r[2], r[3]) if (c & 1) else (r[0], r[1])
        bits = (hi << np.uint64(21)) | (lo >> np.uint64(11))
        u = bits.astype(np.float64) * 2.0 ** -53
        # fma(u, w, l): u w is exact in long double for 53-bit u and w, so one rounding after the sum == fma
        w = np.longdouble(upper[c] - lower[c])
        out[:, c] = (u.astype(np.longdouble) * w + np.longdouble(lower[c])).astype(np.float64)
    return out
