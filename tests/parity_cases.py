"""Parity cases at the BASELINE.json configs' own shapes: the CUDA path (through the C ABI) against the float64 oracle
and, on a subset of points, both against the extended-precision restatement (oracle/hiprec.py).

Shared by tests/test_gpu_configs.py (asserts) and tools/parity_table.py (profiles/parity_r02.md).  Every case returns
rows  (config, quantity, points, err GPU-vs-oracle, err GPU-vs-hiprec, err oracle-vs-hiprec)  with the error measure

    relerr(a, ref) = max_i |a_i - ref_i| / max(|ref_i|, 1e-12 max_j |ref_j|)      EI, sd, qEI, AKG (positive quantities)
    scalederr      = max_i |a_i - ref_i| / (max|y| + sqrt(sd2))                  mean, kriging.quantile: signed
                     quantities that cross zero, measured against the response scale as the round-1 tests do
    grad_relerr    = per candidate, max over the d components divided by the largest component of that gradient.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

QN01 = -1.2815515655446004  # qnorm(0.1)


def relerr(a, ref, floor_frac=1e-12):
    a = np.asarray(a, dtype=np.longdouble); ref = np.asarray(ref, dtype=np.longdouble)
    ok = ~np.isnan(ref)
    if not ok.any():
        return 0.0
    assert np.array_equal(np.isnan(a), ~ok), "NaN pattern differs"
    den = np.maximum(np.abs(ref[ok]), floor_frac * np.max(np.abs(ref[ok])))
    den = np.where(den == 0, 1.0, den)
    return float(np.max(np.abs(a[ok] - ref[ok]) / den))


def scalederr(a, ref, scale):
    a = np.asarray(a, dtype=np.longdouble); ref = np.asarray(ref, dtype=np.longdouble)
    return float(np.max(np.abs(a - ref)) / scale)


def grad_relerr(g, ref):
    g = np.asarray(g, dtype=np.longdouble); ref = np.asarray(ref, dtype=np.longdouble)
    g = g.reshape(-1, g.shape[-1]); ref = ref.reshape(-1, ref.shape[-1])
    scale = np.max(np.abs(ref), axis=1)
    scale = np.maximum(scale, 1e-12 * np.max(scale))
    scale = np.where(scale == 0, 1.0, scale)
    return float(np.max(np.max(np.abs(g - ref), axis=1) / scale))


def _row(cfg, what, npts, a, o, h):
    return {"config": cfg, "quantity": what, "points": int(npts), "gpu_vs_oracle": a, "gpu_vs_hiprec": o, "oracle_vs_hiprec": h}


def _oracle(km):
    from conftest import to_oracle
    return to_oracle(km)


def ei_family(cfg, N, n_hi, n_grad_hi, full_grid=False, seed_rank=7):
    """mean, sd, EI, EI.grad, kriging.quantile(.grad) at a config's model: N candidates vs the oracle, the first
    n_hi (n_grad_hi for gradients) also vs long double."""
    from diceoptim_b200 import synth, predict, EI, EI_grad, kriging_quantile, kriging_quantile_grad
    from oracle import dicekriging as dk, diceoptim as do, hiprec as hp
    km = synth.config_model(cfg)
    ko = _oracle(km)
    x = synth.config_candidates(cfg) if full_grid else synth.config_candidates(cfg, N=N, rank=seed_rank)
    N = x.shape[0]
    rows = []
    got = predict(km, x)
    ref = dk.predict_km(ko, x, light_return=True)
    hi_idx = np.linspace(0, N - 1, n_hi).astype(int)
    hi = hp.predict_km(ko, x[hi_idx])
    scale = float(np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2))
    rows.append(_row(cfg, "mean (/ response scale)", N, scalederr(got["mean"], ref["mean"], scale),
                     scalederr(got["mean"][hi_idx], hi["mean"], scale), scalederr(ref["mean"][hi_idx], hi["mean"], scale)))
    rows.append(_row(cfg, "sd", N, relerr(got["sd"], ref["sd"]), relerr(got["sd"][hi_idx], hi["sd"]), relerr(ref["sd"][hi_idx], hi["sd"])))
    ei = EI(x, km); ei_ref = do.EI(x, ko); ei_hi = hp.EI(x[hi_idx], ko, pred=hi)
    rows.append(_row(cfg, "EI", N, relerr(ei, ei_ref), relerr(ei[hi_idx], ei_hi), relerr(ei_ref[hi_idx], ei_hi)))
    q = kriging_quantile(x, km, beta=0.1); q_ref = do.kriging_quantile(x, ko, beta=0.1)
    q_hi = hp.kriging_quantile(x[hi_idx], ko, QN01, pred=hi)
    rows.append(_row(cfg, "kriging.quantile (/ response scale)", N, scalederr(q, q_ref, scale), scalederr(q[hi_idx], q_hi, scale),
                     scalederr(q_ref[hi_idx], q_hi, scale)))
    g = EI_grad(x, km)
    g_ref = np.stack([do.EI_grad(x[i], ko) for i in range(N)])
    gi = hi_idx[:n_grad_hi]
    g_hi = np.stack([hp.EI_grad(x[i], ko) for i in gi])
    rows.append(_row(cfg, "EI.grad", N, grad_relerr(g, g_ref), grad_relerr(g[gi], g_hi), grad_relerr(g_ref[gi], g_hi)))
    Ng = min(N, 512)
    qg = kriging_quantile_grad(x[:Ng], km, beta=0.1)
    qg_ref = np.stack([do.kriging_quantile_grad(x[i], ko, beta=0.1) for i in range(Ng)])
    gi = gi[gi < Ng]
    qg_hi = np.stack([hp.kriging_quantile_grad(x[i], ko, QN01) for i in gi])
    rows.append(_row(cfg, "kriging.quantile.grad", Ng, grad_relerr(qg, qg_ref), grad_relerr(qg[gi], qg_hi), grad_relerr(qg_ref[gi], qg_hi)))
    argmax_equal = int(np.nanargmax(ei)) == do.which_max(ei_ref)
    km.release()
    return rows, {"argmax_equal": bool(argmax_equal)}


# ---- qEI at cfg3: the oracle's CDFs are pure-Python quadrature (0.3 s per batch) -> worker processes ----
_W = {}


def _qei_worker_init(cfg):
    from diceoptim_b200 import synth
    _W["km"] = _oracle(synth.config_model(cfg))


def _qei_worker(args):
    from oracle import diceoptim as do
    xb, plugin, want_grad = args
    v = do.qEI(xb, _W["km"], plugin=plugin, fastCompute=False)
    g = do.qEI_grad(xb, _W["km"], plugin=plugin, fastCompute=False) if want_grad else None
    return v, g


def qei_family(cfg="cfg3_rosenbrock4_n500", B=256, q=4, workers=None):
    """qEI and qEI.grad for B batches of q points at cfg3's model against the oracle's exact formulas."""
    import multiprocessing as mpx
    from diceoptim_b200 import synth, qEI, qEI_grad
    km = synth.config_model(cfg)
    rng = np.random.default_rng([synth.CONFIGS[cfg]["cand_seed"], 99])
    xb = rng.random((B, q, km.d))
    plugin = float(np.median(km.y))
    got = qEI(xb, km, plugin=plugin)
    gg = qEI_grad(xb, km, plugin=plugin)
    workers = workers or min(os.cpu_count() or 1, 16)
    ctx = mpx.get_context("spawn")      # the parent holds a CUDA context: no fork
    with ctx.Pool(workers, initializer=_qei_worker_init, initargs=(cfg,)) as pool:
        res = pool.map(_qei_worker, [(xb[b], plugin, True) for b in range(B)], chunksize=max(1, B // (4 * workers)))
    ref = np.array([r[0] for r in res]); gref = np.stack([r[1] for r in res])
    km.release()
    scale = float(np.max(np.abs(ref)))
    rows = [_row(cfg, f"qEI (q={q})", B, relerr(got, ref), None, None),
            _row(cfg, f"qEI.grad (q={q})", B, grad_relerr(gg.reshape(B, -1), gref.reshape(B, -1)), None, None)]
    return rows, {"qei_max": scale}


def oracle_predX(ko, type="UK", chunk=500):
    """predict_km(model, model@X) in column chunks (the one-shot call needs an n x n x d temporary)."""
    from oracle import dicekriging as dk
    means, V = [], []
    for i in range(0, ko.X.shape[0], chunk):
        pr = dk.predict_km(ko, ko.X[i:i + chunk], type=type)
        means.append(pr["mean"]); V.append(pr["Tinv.c"])
    return {"mean": np.concatenate(means), "Tinv.c": np.hstack(V)}


def akg_family(cfg="cfg5_noisy20_n5000", noise_kind="noise_var", N=256, n_hi=32, n=None):
    """AKG and kriging.quantile at cfg5's model (noise.var or homogeneous nugget) against the oracle."""
    from diceoptim_b200 import synth, AKG, kriging_quantile, predict
    from oracle import dicekriging as dk, diceoptim as do, hiprec as hp
    km = synth.config_model(cfg, n=n, noise_kind=noise_kind)
    ko = _oracle(km)
    tau2 = synth.CONFIGS[cfg]["noise"]
    x = synth.config_candidates(cfg, N=N, rank=11)
    tag = f"{cfg} ({'noise.var' if noise_kind == 'noise_var' else 'nugget'})"
    rows = []
    got = predict(km, x); ref = dk.predict_km(ko, x, light_return=True)
    hi_idx = np.linspace(0, N - 1, n_hi).astype(int)
    hi = hp.predict_km(ko, x[hi_idx])
    scale = float(np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2))
    rows.append(_row(tag, "mean (/ response scale)", N, scalederr(got["mean"], ref["mean"], scale),
                     scalederr(got["mean"][hi_idx], hi["mean"], scale), scalederr(ref["mean"][hi_idx], hi["mean"], scale)))
    rows.append(_row(tag, "sd", N, relerr(got["sd"], ref["sd"]), relerr(got["sd"][hi_idx], hi["sd"]), relerr(ref["sd"][hi_idx], hi["sd"])))
    qv = kriging_quantile(x, km, beta=0.1); q_ref = do.kriging_quantile(x, ko, beta=0.1)
    q_hi = hp.kriging_quantile(x[hi_idx], ko, QN01, pred=hi)
    rows.append(_row(tag, "kriging.quantile (/ response scale)", N, scalederr(qv, q_ref, scale), scalederr(qv[hi_idx], q_hi, scale),
                     scalederr(q_ref[hi_idx], q_hi, scale)))
    a = AKG(x, km, new_noise_var=tau2)
    predX = oracle_predX(ko)
    a_ref = np.array([do.AKG(x[i], ko, new_noise_var=tau2, predX=predX) for i in range(N)])
    rows.append(_row(tag, "AKG", N, relerr(a, a_ref), None, None))
    rows.append(_row(tag, "AKG (/ response scale)", N, scalederr(a, a_ref, scale), None, None))
    km.release()
    return rows, {"akg_max": float(np.max(a_ref)), "akg_min": float(np.min(a_ref)), "akg": a, "akg_ref": a_ref}


def akg_hiprec(cfg="cfg5_noisy20_n5000", n=300, N=32):
    """AKG at a reduced design size where the reference's own formula (V.X' V.x, an n^3 prediction at the design)
    is affordable in long double: GPU and oracle against it."""
    from diceoptim_b200 import synth, AKG
    from oracle import diceoptim as do, hiprec as hp
    km = synth.config_model(cfg, n=n)
    ko = _oracle(km)
    tau2 = synth.CONFIGS[cfg]["noise"]
    x = synth.config_candidates(cfg, N=N, rank=12)
    a = AKG(x, km, new_noise_var=tau2)
    predX = oracle_predX(ko)
    a_ref = np.array([do.AKG(x[i], ko, new_noise_var=tau2, predX=predX) for i in range(N)])
    predX_hi = hp.predict_km(ko, ko.X)
    a_hi = np.array([hp.AKG(x[i], ko, tau2, predX=predX_hi) for i in range(N)])
    km.release()
    scale = float(np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2))
    tag = f"{cfg} at n={n} (noise.var)"
    return [_row(tag, "AKG", N, relerr(a, a_ref), relerr(a, a_hi), relerr(a_ref, a_hi)),
            _row(tag, "AKG (/ response scale)", N, scalederr(a, a_ref, scale), scalederr(a, a_hi, scale), scalederr(a_ref, a_hi, scale))], \
        {"akg_max": float(np.max(a_ref))}
