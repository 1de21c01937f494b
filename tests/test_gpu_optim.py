"""max_crit / max_EI / max_qEI with the batched polish (all restarts in one launch per iteration) against the
one-restart-at-a-time scipy L-BFGS-B path that mirrors R's optim(); optimisers are compared on the value reached
(1e-6 relative), not on the path."""
import numpy as np
import pytest

from conftest import small_model

pytestmark = pytest.mark.gpu


def _projected_grad_norm(g, x, lo, up):
    g = np.array(g, float)
    g[(x <= lo) & (g < 0)] = 0.0      # maximisation: at the lower bound only a positive slope can be followed
    g[(x >= up) & (g > 0)] = 0.0
    return float(np.max(np.abs(g)))


def test_max_crit_batched_matches_sequential_polish(gpu_lib):
    from diceoptim_b200 import EI, EI_grad, max_crit, synth
    km = synth.config_model("cfg1_branin2_n20")
    lo, up = np.zeros(2), np.ones(2)
    a = max_crit(km, lower=lo, upper=up, control={"restarts": 6, "n.ei": 1000}, seed=5)
    b = max_crit(km, lower=lo, upper=up, control={"restarts": 6, "n.ei": 1000, "optimizer": "scipy"}, seed=5)
    assert np.all(a["par"] >= lo) and np.all(a["par"] <= up)
    assert a["value"] >= b["value"] * (1 - 1e-6)
    assert abs(EI(a["par"], km)[0] - a["value"]) <= 1e-12 * abs(a["value"])
    assert _projected_grad_norm(EI_grad(a["par"], km), a["par"], lo, up) <= 1e-4 * a["value"] + 1e-10
    no_polish = max_crit(km, lower=lo, upper=up, control={"restarts": 6, "n.ei": 1000, "BFGS": False}, seed=5)
    assert a["value"] >= no_polish["value"]
    km.release()


def test_max_ei_screen_then_polish(gpu_lib):
    from diceoptim_b200 import EI, max_EI, screen_EI
    km = small_model("matern5_2", "const", n=60, d=3, seed=21)
    lo, up = np.zeros(3), np.ones(3)
    res = max_EI(km, lower=lo, upper=up, control={"n.screen": 20000, "pop.size": 16}, seed=3)
    x = lo + np.random.default_rng(3).random((20000, 3)) * (up - lo)
    _, v = screen_EI(x, km)
    assert res["value"] >= v[0] and np.all(res["par"] >= lo) and np.all(res["par"] <= up)
    assert abs(EI(res["par"], km)[0] - res["value"]) <= 1e-12 * abs(res["value"])
    # maximisation of the response (the reference keeps plugin = min(y) there, R/max_EI.R:187)
    rmax = max_EI(km, lower=lo, upper=up, minimization=False, control={"n.screen": 5000, "pop.size": 8}, seed=4)
    assert abs(EI(rmax["par"], km, plugin=float(np.min(km.y)), minimization=False)[0] - rmax["value"]) <= 1e-12 * abs(rmax["value"])
    km.release()


def test_max_qei_batched_matches_sequential(gpu_lib):
    from diceoptim_b200 import max_qEI, qEI
    km = small_model("matern5_2", "const", n=50, d=2, seed=8)
    lo, up = np.zeros(2), np.ones(2)
    par0 = np.random.default_rng(2).random((5, 2, 2))
    a = max_qEI(km, 2, lo, up, optimcontrol={"parinit": par0, "maxit": 60})
    b = max_qEI(km, 2, lo, up, optimcontrol={"parinit": par0, "maxit": 60, "optimizer": "scipy"})
    assert a["par"].shape == (2, 2) and np.all(a["par"] >= lo) and np.all(a["par"] <= up)
    assert a["value"] >= b["value"] * (1 - 1e-4), (a["value"], b["value"])
    # the device polish reports qEI at x itself; qEI() rounds x to 8 digits first (R/qEI.R:126-128)
    assert abs(qEI(a["par"], km) - a["value"]) <= 1e-6 * abs(a["value"]) + 1e-12
    start_vals = qEI(par0, km)
    assert a["value"] >= np.max(start_vals) * (1 - 1e-6)
    km.release()


def test_max_qei_constant_liar(gpu_lib):
    """max_qEI(crit = "CL"), R/max_qEI.R:283-318: every round's point maximises EI of the model that already holds the
    earlier points with the lie as response; values are the lies."""
    from diceoptim_b200 import EI, max_EI, max_qEI
    from diceoptim_b200.update import update_km
    km = small_model("matern5_2", "const", n=40, d=2, seed=6)
    lo, up = np.zeros(2), np.ones(2)
    ctl = {"n.screen": 20000, "pop.size": 12}
    res = max_qEI(km, 3, lo, up, crit="CL", optimcontrol=ctl, seed=11)
    assert res["par"].shape == (3, 2) and res["value"].shape == (3, 1)
    assert np.all(res["value"] == np.min(km.y))
    assert np.all(res["par"] >= lo) and np.all(res["par"] <= up)
    cur = km
    for s in range(3):
        best = max_EI(cur, lower=lo, upper=up, control={"n.screen": 50000, "pop.size": 24}, seed=100 + s)
        got = EI(res["par"][s], cur, plugin=float(np.min(cur.y)))[0]
        assert got >= best["value"] * (1 - 1e-3), (s, got, best["value"])      # same optimum from an independent run
        nxt = update_km(cur, res["par"][s], float(np.min(km.y)))
        if cur is not km:
            cur.release()
        cur = nxt
    cur.release()
    kb = max_qEI(km, 2, lo, up, crit="CL", optimcontrol={**ctl, "L": "mean"}, seed=3)
    assert np.all(np.isfinite(kb["value"])) and not np.all(kb["value"] == np.min(km.y))
    with pytest.raises(ValueError):
        max_qEI(km, 2, lo, up, crit="CL", optimcontrol={"L": "median"})
    km.release()


def test_noisy_criteria_maximisers(gpu_lib):
    """max_AEI / max_AKG / max_EQI / min_quantile (R/max_AEI.R, R/max_AKG.R, R/max_EQI.R, R/min_quantile.R): the
    returned value is the criterion at the returned point, not below the best of an independent random screen, and
    the point is stationary up to the box."""
    from diceoptim_b200 import (AEI, AKG, EQI, kriging_quantile, max_AEI, max_AKG, max_EQI, min_quantile)
    km = small_model("matern5_2", "const", n=40, d=2, seed=12, noise_var=0.05)
    lo, up = np.zeros(2), np.ones(2)
    xs = np.random.default_rng(9).random((4000, 2))
    ctl = {"n.screen": 4000, "pop.size": 12}
    cases = [(max_AEI, lambda X: AEI(X, km, new_noise_var=0.05), dict(new_noise_var=0.05), True),
             (max_AKG, lambda X: AKG(X, km, new_noise_var=0.05), dict(new_noise_var=0.05), True),
             (max_EQI, lambda X: EQI(X, km, new_noise_var=0.05), dict(new_noise_var=0.05), True),
             (min_quantile, lambda X: kriging_quantile(X, km, beta=0.1), dict(beta=0.1), False)]
    for fn, crit, kw, maximise in cases:
        res = fn(km, lower=lo, upper=up, control=ctl, seed=1, **kw)
        assert res["par"].shape == (1, 2) and np.all(res["par"] >= lo) and np.all(res["par"] <= up)
        at = float(np.atleast_1d(crit(res["par"]))[0])
        assert abs(at - res["value"]) <= 1e-9 * abs(at) + 1e-14, (fn.__name__, at, res["value"])
        scr = crit(xs)
        if maximise:
            assert res["value"] >= np.nanmax(scr) * (1 - 1e-9), fn.__name__
        else:
            assert res["value"] <= np.nanmin(scr) + 1e-9 * abs(np.nanmin(scr)), fn.__name__
    km.release()


def test_device_polish_matches_host_batched_lbfgs(gpu_lib):
    """dk_polish (optimiser state on the device) against the host-side batched L-BFGS (optim.py) from the same starts:
    same method, so the optima agree to optimiser tolerance; both must sit on a stationary point of the box problem."""
    from diceoptim_b200 import EI, EI_grad, synth
    from diceoptim_b200.api import _ei_value_and_grad
    from diceoptim_b200.optim import batched_lbfgsb, device_polish
    km = small_model("matern5_2", "const", n=80, d=4, seed=33)
    lo, up = np.zeros(4), np.ones(4)
    x0 = np.random.default_rng(4).random((24, 4))
    pl = float(np.min(km.y))
    dev = device_polish(km, "EI", x0, lo, up, plugin=pl, maxit=80)

    def fg(X):
        ei, gr = _ei_value_and_grad(X, km, pl, "UK", True, False, 0)
        return -ei, -np.asarray(gr)
    host = batched_lbfgsb(fg, x0, lo, up, maxit=80)
    start = EI(x0, km)
    assert np.all(dev["f"] >= start * (1 - 1e-12) - 1e-300)                 # never worse than the start
    assert np.all(dev["x"] >= lo) and np.all(dev["x"] <= up)
    assert np.allclose(EI(dev["x"], km), dev["f"], rtol=1e-12, atol=1e-300)   # the reported value is the criterion there
    # same local optima for most starts (asynchronous vs synchronous line searches may pick different basins rarely)
    close = np.abs(dev["f"] - (-host["f"])) <= 1e-6 * np.maximum(np.abs(host["f"]), 1e-12)
    assert close.mean() >= 0.75, (dev["f"], -host["f"])
    assert dev["f"].max() >= (-host["f"]).max() * (1 - 1e-6)
    g = EI_grad(dev["x"], km)
    best = int(np.argmax(dev["f"]))
    assert _projected_grad_norm(g[best], dev["x"][best], lo, up) <= 1e-3 * dev["f"][best] + 1e-10    # stopped by factr, as optim
    km.release()


def test_device_polish_quantile_and_qei(gpu_lib):
    from diceoptim_b200 import kriging_quantile, qEI
    from diceoptim_b200.optim import device_polish
    km = small_model("matern5_2", "const", n=60, d=2, seed=34, noise_var=0.02)
    lo, up = np.zeros(2), np.ones(2)
    x0 = np.random.default_rng(5).random((8, 2))
    res = device_polish(km, "quantile", x0, lo, up, qnorm_beta=-1.2815515655446004, maxit=60)
    q0 = kriging_quantile(x0, km, beta=0.1)
    assert np.all(res["f"] <= q0 + 1e-12)                                   # a minimisation
    assert np.allclose(kriging_quantile(res["x"], km, beta=0.1), res["f"], rtol=1e-12)
    xb = np.random.default_rng(6).random((5, 3, 2))
    pl = float(np.median(km.y))
    r2 = device_polish(km, "qEI", xb.reshape(5, 6), lo, up, q=3, plugin=pl, maxit=40)
    v0 = qEI(xb, km, plugin=pl)
    assert np.all(r2["f"] >= v0 * (1 - 1e-6))
    # dk_polish takes qEI at x itself (one pass with the gradient), dk_qei at round(x, 8) as R/qEI.R:126-128 does
    assert np.allclose(qEI(r2["x"].reshape(5, 3, 2), km, plugin=pl), r2["f"], rtol=1e-6, atol=1e-9)
    # q = 5: value and gradient from the general-q kernels (lattice CDFs)
    xb5 = np.random.default_rng(7).random((3, 5, 2))
    r5 = device_polish(km, "qEI", xb5.reshape(3, 10), lo, up, q=5, plugin=pl, maxit=15)
    v5 = qEI(xb5, km, plugin=pl)
    assert np.all(r5["f"] >= v5 * (1 - 1e-6)) and np.any(r5["f"] > v5 * 1.01)
    assert np.allclose(qEI(r5["x"].reshape(3, 5, 2), km, plugin=pl), r5["f"], rtol=1e-6, atol=1e-9)
    km.release()


def test_sample_from_ei_on_device_follows_the_weights(gpu_lib):
    """dk_sample_from_ei: draws without replacement with probability proportional to the EI weights.  Statistical check of
    the first draw against the weights (chi-square over 8 bins of a small population), distinctness, determinism by seed."""
    import ctypes as C
    from diceoptim_b200 import _lib
    km = small_model("matern5_2", "const", n=40, d=2, seed=35)
    h = km.handle(0)
    N = 64
    x = np.asfortranarray(np.random.default_rng(7).random((N, 2)))
    T = float(np.min(km.y))
    w = np.empty(N)
    _lib.check(gpu_lib.dk_sample_weights(h.ptr, C.c_void_p(x.ctypes.data), N, T, 1, C.c_void_p(w.ctypes.data)), "weights")
    assert np.count_nonzero(w) >= 8
    p = w / w.sum()
    trials = 4000
    first = np.zeros(N)
    idx = np.empty(5, np.int64)
    for s in range(trials):
        _lib.check(gpu_lib.dk_sample_from_ei(h.ptr, C.c_void_p(x.ctypes.data), N, T, 1, 5, s + 1, C.c_void_p(idx.ctypes.data)), "draw")
        assert len(set(idx.tolist())) == 5 and idx.min() >= 0
        first[idx[0]] += 1
    order = np.argsort(-p)
    bins = np.array_split(order, 8)
    obs = np.array([first[b].sum() for b in bins]); exp = np.array([p[b].sum() for b in bins]) * trials
    keep = exp > 5
    chi2 = float(np.sum((obs[keep] - exp[keep]) ** 2 / exp[keep]))
    assert chi2 < 30.0, (chi2, obs, exp)                                   # 7 d.o.f.: P(chi2 > 30) ~ 1e-4
    a = np.empty(5, np.int64); b = np.empty(5, np.int64)
    gpu_lib.dk_sample_from_ei(h.ptr, C.c_void_p(x.ctypes.data), N, T, 1, 5, 99, C.c_void_p(a.ctypes.data))
    gpu_lib.dk_sample_from_ei(h.ptr, C.c_void_p(x.ctypes.data), N, T, 1, 5, 99, C.c_void_p(b.ctypes.data))
    assert np.array_equal(a, b)
    km.release()


def test_model_append_in_place_matches_rebuild(gpu_lib):
    """dk_model_append (in-place growth of the device handle: last block rows of the tiled factor only) against a handle
    created from scratch for the grown model: forward path (EI), gradient path (backward tiles rebuilt lazily), AKG
    constants, qEI's duplicate rule on the appended row; across a block-row boundary and a capacity growth."""
    from diceoptim_b200 import AKG, EI, EI_grad, qEI, synth
    from diceoptim_b200.update import update_km
    X = synth.lhs(250, 3, 5); y = np.sin(3 * X.sum(1))
    km = synth.make_km(X, y, "matern5_2", 0.5, float(np.var(y)), trend="linear", noise_var=0.01)
    x = np.random.default_rng(3).random((300, 3))
    EI(x, km); EI_grad(x[:20], km); AKG(x[:20], km, new_noise_var=0.01)      # build the lazily made pieces first
    pts = np.random.default_rng(4).random((12, 3))
    cur = km
    for i, xn in enumerate(pts):                                            # 250 -> 262 rows: crosses 256, grows capacity
        cur = update_km(cur, xn, float(np.cos(i)), new_noise_var=0.0)
        assert cur._handles and not km._handles or i > 0
    assert gpu_lib.dk_model_n(cur.handle(0).ptr) == 262
    fresh = update_km(km, pts[0], 1.0, reuse_handles=False)
    for i, xn in enumerate(pts[1:], 1):
        fresh = update_km(fresh, xn, float(np.cos(i)), reuse_handles=False)
    a = EI(x, cur); b = EI(x, fresh)
    assert np.all(np.abs(a - b) <= 1e-9 * np.abs(b) + 1e-12 * b.max())
    ga = EI_grad(x[:40], cur); gb = EI_grad(x[:40], fresh)
    assert np.max(np.abs(ga - gb)) <= 1e-9 * np.max(np.abs(gb))
    ka = AKG(x[:40], cur, new_noise_var=0.01); kb = AKG(x[:40], fresh, new_noise_var=0.01)
    assert np.all(np.abs(ka - kb) <= 1e-7 * np.abs(kb) + 1e-11)
    xb = np.stack([np.vstack([pts[3], x[0], x[1]]), x[2:5]])               # batch 0 holds an appended design row
    qa = qEI(xb, cur, plugin=float(np.median(y))); qb = qEI(xb, fresh, plugin=float(np.median(y)))
    assert np.all(np.abs(qa - qb) <= 1e-9 * np.abs(qb) + 1e-12)
    cur.release(); fresh.release()
