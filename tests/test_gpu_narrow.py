"""Narrow tiles (16 candidates per CTA, csrc/dk_narrow.cuh) against the wide path and the oracle.  The solved panel is
bit-identical by construction (same covariance values, same accumulation order), so the mean must agree bit for bit;
sd, EI and the gradients differ only through the grouping of the row reductions (8 partial sums of v^2 instead of 4, 16
row slices of the gradient contraction instead of 4): 1e-14 of sd2 in the variance.  EI and the gradients are functions
of that sd: in the far tail (|xcr| ~ 30) or on an ill-conditioned gauss design a last-bit change of sd^2 moves them by
1e-10 .. 1e-9 relative, on either path; they are compared here to 1e-8 and held to the oracle in the other test files
(which run on narrow tiles whenever N <= 4736)."""
import ctypes as C
import time

import numpy as np
import pytest

from conftest import small_model, to_oracle

pytestmark = pytest.mark.gpu
WIDE, NARROW = 1, 2


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _both(gpu_lib, km, x, type=1):
    from diceoptim_b200 import _lib
    h = km.handle(0)
    N, d = x.shape
    x = np.asfortranarray(x)
    pl = float(np.median(km.y))      # a reachable target: EI of order 0.1 everywhere, not 1e-200 tails
    out = {}
    for mode in (WIDE, NARROW):
        _lib.check(gpu_lib.dk_model_set_tile_mode(h.ptr, mode), "tile mode")
        ei = np.empty(N); mean = np.empty(N); sd = np.empty(N); g = np.empty((N, d), order="F"); e2 = np.empty(N)
        q = np.empty(N); gq = np.empty((N, d), order="F")
        _lib.check(gpu_lib.dk_ei(h.ptr, _vp(x), N, pl, type, 1, 0, _vp(ei), _vp(mean), _vp(sd)), "dk_ei")
        _lib.check(gpu_lib.dk_ei_grad(h.ptr, _vp(x), N, pl, type, 1, 0, _vp(g), _vp(e2)), "dk_ei_grad")
        _lib.check(gpu_lib.dk_quantile(h.ptr, _vp(x), N, -0.84, type, _vp(q), _vp(gq)), "dk_quantile")
        out[mode] = dict(ei=ei, mean=mean, sd=sd, g=g, e2=e2, q=q, gq=gq)
    gpu_lib.dk_model_set_tile_mode(h.ptr, 0)
    return out


@pytest.mark.parametrize("name,trend,kw", [("matern5_2", "const", {}), ("gauss", "linear", {}), ("exp", "const", {"nugget": 0.02}),
                                           ("powexp", "const", {}), ("matern3_2", "linear", {"noise_var": 0.01})])
@pytest.mark.parametrize("N", [1, 15, 16, 17, 100])
def test_narrow_equals_wide(gpu_lib, name, trend, kw, N):
    km = small_model(name, trend, n=300, d=4, seed=41, **kw)
    x = np.random.default_rng(N).random((N, 4))
    x[0] = km.X[7]                                   # a design point: sd -> 0 guards
    o = _both(gpu_lib, km, x)
    w, nr = o[WIDE], o[NARROW]
    assert np.array_equal(w["mean"], nr["mean"])    # same panel, same epilogue
    tot = km.covariance.sd2 + (km.covariance.nugget if km.covariance.nugget_flag else 0.0)
    assert np.max(np.abs(w["sd"] ** 2 - nr["sd"] ** 2)) <= 1e-14 * tot
    for key in ("ei", "e2", "q"):
        assert np.array_equal(np.isnan(w[key]), np.isnan(nr[key]))
        ok = ~np.isnan(w[key])
        assert np.all(np.abs(w[key][ok] - nr[key][ok]) <= 1e-8 * np.abs(w[key][ok]) + 1e-12 * np.max(np.abs(w[key][ok]) + 1e-300)), key
    for key in ("g", "gq"):
        scale = np.max(np.abs(w[key]), axis=1, keepdims=True) + 1e-300
        assert np.max(np.abs(w[key] - nr[key]) / scale) <= 1e-8, key
    km.release()


def test_narrow_vs_oracle_multi_block_and_timing(gpu_lib):
    """n = 2000 (16 block rows): a 16-start batch, values and gradients against the oracle, and the latency the mode is for."""
    from diceoptim_b200 import synth, _lib
    from oracle import diceoptim as do
    km = synth.config_model("cfg4_sphere10_n2000")
    ko = to_oracle(km)
    x = synth.config_candidates("cfg4_sphere10_n2000", N=16, rank=5)
    o = _both(gpu_lib, km, x)
    plm = float(np.median(km.y))                        # _both's plugin
    ref = do.EI(x, ko, plugin=plm)
    for mode in (WIDE, NARROW):
        assert np.all(np.abs(o[mode]["ei"] - ref) <= 1e-10 * np.abs(ref) + 1e-12 * ref.max())
    for j in (0, 7, 15):
        gr = do.EI_grad(x[j], ko, plugin=plm)
        assert np.max(np.abs(o[NARROW]["g"][j] - gr)) <= 1e-10 * np.max(np.abs(gr)) + 1e-14
    # latency of one 16-candidate call through the host entry points (H2D + kernel + D2H), best of 5
    h = km.handle(0)
    xf = np.asfortranarray(x); ei = np.empty(16); g = np.empty((16, km.d), order="F")
    pl = float(np.min(km.y))
    times = {}
    for mode in (WIDE, NARROW):
        gpu_lib.dk_model_set_tile_mode(h.ptr, mode)
        best_f = best_g = 1e9
        for _ in range(5):
            t0 = time.perf_counter(); _lib.check(gpu_lib.dk_ei(h.ptr, _vp(xf), 16, pl, 1, 1, 0, _vp(ei), None, None), "ei")
            best_f = min(best_f, time.perf_counter() - t0)
            t0 = time.perf_counter(); _lib.check(gpu_lib.dk_ei_grad(h.ptr, _vp(xf), 16, pl, 1, 1, 0, _vp(g), _vp(ei)), "grad")
            best_g = min(best_g, time.perf_counter() - t0)
        times[mode] = (best_f * 1e3, best_g * 1e3)
    gpu_lib.dk_model_set_tile_mode(h.ptr, 0)
    print(f"\\n16-start batch at n=2000: wide {times[WIDE][0]:.2f} / {times[WIDE][1]:.2f} ms, narrow {times[NARROW][0]:.2f} / {times[NARROW][1]:.2f} ms (EI / EI+grad)")
    assert times[NARROW][0] < 1.0 and times[NARROW][1] < 2.5      # VERDICT r01 targets for this shape
    km.release()
