"""The single-block kernel for small designs (n + p + 1 <= 256, csrc/dk_small.cuh) against the oracle and against the
blocked kernel: it serves every batch of such a model that is too large for the narrow tiles (N > 4736), i.e. the grid /
random-search screens of BASELINE cfg1 and cfg2."""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_close_adjudicated, small_model, to_oracle

pytestmark = pytest.mark.gpu
N_BIG = 6000        # 94 wide tiles: neither narrow nor (for small designs) blocked


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


ULPS = 16 * 2.220446049250313e-16      # allowance for the covariance vector: exp (1 ulp) and, per dimension, the scaled
                                        # distance |X w - x w| (two rounded products whose difference loses up to |X w| / a)


def _within(got, ref, tol, hi_of, sens_of, what):
    """|got - ref| <= tol against the float64 oracle; where that fails (ill-conditioned gauss designs, on which the oracle's
    own rounding is of that size) the long-double restatement decides at the 48 worst entries: the GPU is then within tol
    of it, or no further from it than 4x the oracle is there, or than a 16-ulp relative error of the covariance vector
    moves the exact value (oracle/hiprec.predict_sensitivity: no float64 evaluation of c avoids a few ulp)."""
    err = np.abs(got - ref)
    bad = np.flatnonzero(~(err <= tol))
    if bad.size == 0:
        return
    bad = bad[np.argsort(-err[bad])[:48]]
    hi = np.asarray(hi_of(bad), dtype=np.longdouble)
    g = np.abs(got[bad] - hi); o = np.abs(ref[bad] - hi)
    allowed = np.maximum(tol, np.maximum(4.0 * o, ULPS * np.asarray(sens_of(bad), dtype=np.longdouble)))
    assert np.all(g <= allowed), (what, "GPU vs long double", float(g.max()), "oracle vs long double", float(o.max()),
                                  "tol", tol, "worst excess", float(np.max(g / allowed)))


def _ei(gpu_lib, km, x, mode, plugin, type=1, minimization=1):
    from diceoptim_b200 import _lib
    h = km.handle(0)
    N = x.shape[0]
    x = np.asfortranarray(x)
    ei = np.empty(N); mean = np.empty(N); sd = np.empty(N)
    _lib.check(gpu_lib.dk_model_set_tile_mode(h.ptr, mode), "tile mode")
    _lib.check(gpu_lib.dk_ei(h.ptr, _vp(x), N, plugin, type, minimization, 0, _vp(ei), _vp(mean), _vp(sd)), "dk_ei")
    gpu_lib.dk_model_set_tile_mode(h.ptr, 0)
    return ei, mean, sd


@pytest.mark.parametrize("n", [1, 20, 100, 126, 127, 200, 253])
@pytest.mark.parametrize("name,trend,kw", [("matern5_2", "const", {}), ("gauss", "linear", {}), ("exp", "const", {"nugget": 0.02}),
                                           ("powexp", "linear", {"noise_var": 0.01})])
def test_small_kernel_matches_oracle_and_blocked_kernel(gpu_lib, n, name, trend, kw):
    d = 3
    if trend == "linear" and n < d + 2:
        pytest.skip("fewer design points than trend columns")
    if n + (d + 1 if trend == "linear" else 1) + 1 > 256:
        pytest.skip("not a small design")
    from oracle import dicekriging as dk, diceoptim as do, hiprec as hp
    km = small_model(name, trend, n=n, d=d, seed=90 + n, **kw)
    ko = to_oracle(km)
    uses_small = bool(gpu_lib.dk_model_small_kernel(km.handle(0).ptr))
    if name != "gauss":
        assert uses_small          # the ill-conditioned gauss designs may be refused by the accuracy probe, nothing else
    x = np.random.default_rng(n).random((N_BIG, d))
    x[5] = km.X[0]                                                   # a design point
    pl = float(np.median(km.y))
    for type_id, type in ((1, "UK"), (0, "SK")):
        ei, mean, sd = _ei(gpu_lib, km, x, 0, pl, type=type_id)      # AUTO -> the small kernel
        eb, mb, sb = _ei(gpu_lib, km, x, 1, pl, type=type_id)        # WIDE -> the blocked kernel
        if not uses_small:
            # refused by the accuracy probe (gauss, n = 200: sum |w_i c_i| = 2.6e6, no float64 evaluation reaches 1e-10):
            # AUTO is then the blocked kernel itself, which the other GPU tests cover
            assert np.array_equal(mean, mb) and np.array_equal(sd, sb) and np.array_equal(ei, eb, equal_nan=True)
            continue
        ref = dk.predict_km(ko, x, type=type, light_return=True)
        scale = np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2)
        tot = km.covariance.sd2 + (km.covariance.nugget if km.covariance.nugget_flag else 0.0)
        hi_pred = lambda bad: hp.predict_km(ko, x[bad], type=type)
        sens = lambda bad: hp.predict_sensitivity(ko, x[bad], type=type)
        _within(mean, ref["mean"], 1e-10 * scale, lambda bad: hi_pred(bad)["mean"], lambda bad: sens(bad)["mean"].ravel(),
                "mean (small kernel)")
        _within(sd ** 2, ref["sd"] ** 2, 1e-10 * tot, lambda bad: np.asarray(hi_pred(bad)["sd"], dtype=np.longdouble) ** 2,
                lambda bad: sens(bad)["sd2"], "sd^2 (small kernel)")
        assert np.max(np.abs(mean - mb)) <= 1e-11 * scale and np.max(np.abs(sd ** 2 - sb ** 2)) <= 1e-11 * tot
        with np.errstate(all="ignore"):
            eref = do.EI(x, ko, plugin=pl, type=type)
        assert np.array_equal(np.isnan(ei), np.isnan(eref))
        ok = np.flatnonzero(~np.isnan(eref))
        floor = 1e-12 * np.max(np.abs(eref[ok]))
        # loose = 1e-7: on the gauss designs the float64 mean itself is off by 1e-11 absolute (oracle 8.6e-12, GPU 3.3e-12
        # against long double at n = 100), which is 1e-8 of an EI of 2e-4; the adjudicated bound above is the real check
        assert_close_adjudicated(ei[ok], eref[ok], lambda bad: hp.EI(x[ok[bad]], ko, plugin=pl, type=type), floor=floor,
                                 loose=1e-7, what="EI (small kernel)")
    km.release()


def test_small_kernel_cfg2_shape_and_other_criteria(gpu_lib):
    """cfg2's own shape (hartman4, gauss, n = 200, d = 4) on 20 000 candidates: EI and which.max against the oracle, the
    quantile epilogue, the host pipeline, and bit-exact repeats."""
    from diceoptim_b200 import synth, EI, kriging_quantile, screen_EI
    from oracle import diceoptim as do
    km = synth.config_model("cfg2_hartman4_n200")
    ko = to_oracle(km)
    assert gpu_lib.dk_model_small_kernel(km.handle(0).ptr) == 1      # cfg2 itself qualifies
    x = synth.config_candidates("cfg2_hartman4_n200", N=20_000, rank=3)
    ei = EI(x, km)
    ref = do.EI(x, ko)
    assert np.all(np.abs(ei - ref) <= 1e-10 * np.abs(ref) + 1e-12 * ref.max())
    idx, val = screen_EI(x, km)
    assert idx[0] == do.which_max(ref) and val[0] == ei[idx[0]]
    q = kriging_quantile(x, km, beta=0.1)
    qref = do.kriging_quantile(x, ko, beta=0.1)
    assert np.max(np.abs(q - qref)) <= 1e-10 * (np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2))
    assert np.array_equal(ei, EI(x, km))
    km.release()


def test_small_kernel_follows_model_append(gpu_lib):
    """dk_model_append re-forms the whole-factor inverse and its shared-memory map (refresh_small_tiles): a model grown in
    place from 100 to 130 rows (one row block -> two) against a handle created for the grown model, both on large batches."""
    from diceoptim_b200 import EI, synth
    from diceoptim_b200.update import update_km
    X = synth.lhs(100, 3, 7); y = np.sin(3 * X.sum(1)) + 0.3 * X[:, 1]
    km = synth.make_km(X, y, "matern5_2", 0.5, float(np.var(y)), trend="const", noise_var=0.01)
    x = np.random.default_rng(8).random((N_BIG, 3))
    EI(x, km)                                                            # the handle exists before it grows
    assert gpu_lib.dk_model_small_kernel(km.handle(0).ptr) == 1
    pts = np.random.default_rng(9).random((30, 3))
    fresh = km                                                           # same updates, handle built from scratch at the end
    for i, xn in enumerate(pts):
        fresh = update_km(fresh, xn, float(np.cos(i)), new_noise_var=0.01, reuse_handles=False)
    cur = km
    for i, xn in enumerate(pts):
        cur = update_km(cur, xn, float(np.cos(i)), new_noise_var=0.01)
    assert gpu_lib.dk_model_n(cur.handle(0).ptr) == 130 and gpu_lib.dk_model_small_kernel(cur.handle(0).ptr) == 1
    pl = float(np.median(y))
    a = EI(x, cur, plugin=pl)
    b, _, _ = _ei(gpu_lib, fresh, x, 1, pl)                               # WIDE: the blocked kernel on a fresh handle
    assert np.all(np.abs(a - b) <= 1e-9 * np.abs(b) + 1e-12 * b.max())
    cur.release(); fresh.release()
