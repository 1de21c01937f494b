"""DiceOptim's acquisition functions with their reference signatures, evaluated by the CUDA engine.

Each function mirrors the R function of the same name (argument names, defaults, error
behaviour); matrices of candidates are evaluated in one fused kernel launch.  Differences from
the reference are limited to what a GPU batch cannot give back: ``envir`` receives the
per-candidate scalars (xcr, xcr.prob, xcr.dens, kriging.sd) but not the n x N matrices ``c`` and
``Tinv.c`` (SURVEY 8(b) "envir contract"); ``qEI(x, envir)`` leaves the finished gradient there for ``qEI_grad(x, envir)``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
from scipy.special import ndtr, ndtri

from . import _lib
from .km import KM, TYPE_IDS

_INV_SQRT_2PI = 0.3989422804014326779399460599343818684758586311649


def _as_candidates(x, d):
    """R/EI.R:93-108: vector -> 1 x d, matrix -> N x d; column-major float64 for the C ABI."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim <= 1:
        x = x.reshape(1, -1)
    if x.ndim != 2 or x.shape[1] != d:
        raise ValueError("x does not have the right size")
    return np.asfortranarray(x)


def _type_id(type):
    if type not in TYPE_IDS:
        raise ValueError(f"type must be 'UK' or 'SK', got {type!r}")
    return TYPE_IDS[type]


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _default_plugin(model: KM, minimization: bool) -> float:
    # R/EI.R:82-88
    return float(np.min(model.y)) if minimization else float(-np.max(model.y))


def predict(model: KM, newdata, type="UK", device=0):
    """predict.km(..., light.return=TRUE): dict(mean=, sd=) (DiceKriging; call site R/EI.R:115)."""
    x = _as_candidates(newdata, model.d)
    N = x.shape[0]
    mean = np.empty(N); sd = np.empty(N)
    h = model.handle(device)
    _lib.check(_lib.load().dk_predict(h.ptr, _vp(x), N, _type_id(type), _vp(mean), _vp(sd)), "dk_predict")
    return {"mean": mean, "sd": sd}


def EI(x, model: KM, plugin=None, type="UK", minimization=True, envir=None, proxy=False, device=0):
    """Expected improvement, R/EI.R:80-145.  Returns one value per row of ``x``."""
    if plugin is None:
        plugin = _default_plugin(model, minimization)
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    ei = np.empty(N)
    want = envir is not None
    mean = np.empty(N) if want else None
    sd = np.empty(N) if want else None
    h = model.handle(device)
    _lib.check(_lib.load().dk_ei(h.ptr, _vp(x), N, float(plugin), _type_id(type), int(bool(minimization)),
                                 int(bool(proxy)), _vp(ei), _vp(mean), _vp(sd)), "dk_ei")
    if want:
        km = mean if minimization else -mean
        if proxy:
            xcr = prob = dens = None
        else:
            with np.errstate(divide="ignore", invalid="ignore"):
                xcr = (plugin - km) / sd
            prob = ndtr(xcr); dens = _INV_SQRT_2PI * np.exp(-0.5 * xcr * xcr)
        envir.update({"xcr": xcr, "xcr.prob": prob, "xcr.dens": dens, "kriging.sd": sd,
                      "kriging.mean": mean, "c": None, "Tinv.c": None})
    return ei


def EI_grad(x, model: KM, plugin=None, type="UK", minimization=True, envir=None, proxy=False, device=0):
    """Analytic gradient of EI, R/EI.grad.R:97-186.  A length-d ``x`` gives a length-d gradient (the
    reference's contract); an N x d matrix gives the N x d matrix of per-row gradients."""
    if plugin is None:
        plugin = _default_plugin(model, minimization)
    single = np.asarray(x).ndim <= 1
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    grad = np.empty((N, model.d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_ei_grad(h.ptr, _vp(x), N, float(plugin), _type_id(type), int(bool(minimization)),
                                      int(bool(proxy)), _vp(grad), None), "dk_ei_grad")
    return grad[0].copy() if single else grad


def _ei_value_and_grad(x, model: KM, plugin, type, minimization, proxy, device):
    """EI and its gradient for every row of ``x`` from ONE launch of the gradient kernel (dk_ei_grad fills both)."""
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    grad = np.empty((N, model.d), order="F"); ei = np.empty(N)
    h = model.handle(device)
    _lib.check(_lib.load().dk_ei_grad(h.ptr, _vp(x), N, float(plugin), _type_id(type), int(bool(minimization)),
                                      int(bool(proxy)), _vp(grad), _vp(ei)), "dk_ei_grad")
    return ei, grad


def kriging_quantile(x, model: KM, beta=0.1, type="UK", envir=None, device=0):
    """Kriging quantile mk + qnorm(beta) sk, R/kriging.quantile.R:92-115."""
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    q = np.empty(N)
    h = model.handle(device)
    _lib.check(_lib.load().dk_quantile(h.ptr, _vp(x), N, float(ndtri(beta)), _type_id(type), _vp(q), None),
               "dk_quantile")
    return q


def kriging_quantile_grad(x, model: KM, beta=0.1, type="UK", envir=None, device=0):
    """Gradient of the kriging quantile, R/kriging.quantile.grad.R:83-138 (rows batched like EI_grad)."""
    single = np.asarray(x).ndim <= 1
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    q = np.empty(N)
    grad = np.empty((N, model.d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_quantile(h.ptr, _vp(x), N, float(ndtri(beta)), _type_id(type), _vp(q), _vp(grad)),
               "dk_quantile")
    return grad[0].copy() if single else grad


def screen_EI(x, model: KM, plugin=None, type="UK", minimization=True, proxy=False, k=1, device=0):
    """The batched screen of R/max_crit.R:104-110 on the device: EI over the rows of ``x`` then
    which.max (first maximal index, NaN skipped).  Returns (indices[k], values[k]), 0-based."""
    if plugin is None:
        plugin = _default_plugin(model, minimization)
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    idx = np.empty(k, dtype=np.int64); val = np.empty(k)
    h = model.handle(device)
    _lib.check(_lib.load().dk_screen_ei(h.ptr, _vp(x), 0, N, float(plugin), _type_id(type), int(bool(minimization)),
                                        int(bool(proxy)), int(k), _vp(idx), _vp(val), None), "dk_screen_ei")
    return idx, val


def max_crit(model: KM, type="UK", lower=None, upper=None, minimization=True, control=None, proxy=False,
             trcontrol=None, n_cores=1, seed=None, device=0):
    """R/max_crit.R:64-130: per restart, N uniform candidates -> EI screen on the GPU -> which.max ->
    optional L-BFGS-B polish with EI / EI.grad; best restart wins.  ``n_cores`` must be 1: the
    reference forks with mclapply and CUDA contexts do not survive fork()."""
    from scipy.optimize import minimize
    if n_cores != 1:
        raise ValueError("n.cores must be 1 with the GPU engine (restarts are batched in-process)")
    control = dict(control or {}); trcontrol = dict(trcontrol or {})
    d = model.d
    lower = np.zeros(d) if lower is None else np.asarray(lower, float)
    upper = np.ones(d) if upper is None else np.asarray(upper, float)
    trcontrol.setdefault("sigma", (upper - lower) / 2)
    trcontrol.setdefault("x0", model.X[int(np.argmin(model.y))])
    trcontrol.setdefault("crit", "EI")
    trcontrol.setdefault("inTR", False)
    if trcontrol["inTR"]:                                   # R/max_crit.R:76-82
        lower = np.maximum(trcontrol["x0"] - trcontrol["sigma"], lower)
        upper = np.minimum(trcontrol["x0"] + trcontrol["sigma"], upper)
        if trcontrol["crit"] != "EI":
            proxy = True
    restarts = control.get("restarts", min(10, d))
    bfgs = control.get("BFGS", True)
    N = int(control.get("n.ei", min(2000, 500 * d)))
    plugin = _default_plugin(model, minimization)
    rng = np.random.default_rng(seed)
    optimizer = control.get("optimizer", "device")
    if optimizer in ("device", "batched"):
        # every restart's screen, then ONE polish of all the restarts' winners together: on the device (dk_polish, the
        # optimiser state never leaves it) or, "batched", with the L-BFGS algebra on the host (optim.py)
        from .optim import batched_lbfgsb, device_polish
        starts = np.empty((restarts, d)); vals = np.empty(restarts)
        for r in range(restarts):
            x = lower + rng.random((N, d)) * (upper - lower)
            idx, val = screen_EI(x, model, plugin=None, type=type, minimization=minimization, proxy=proxy, device=device)
            starts[r], vals[r] = x[idx[0]], val[0]
        if bfgs and optimizer == "device":
            res = device_polish(model, "EI", starts, lower, upper, type=type, minimization=True, proxy=proxy, plugin=plugin,
                                maxit=int(control.get("maxit", 100)), device=device)   # minimization = TRUE: R/max_crit.R:119
            improved = res["f"] >= vals
            starts = np.where(improved[:, None], res["x"], starts); vals = np.where(improved, res["f"], vals)
        elif bfgs:
            def fg(X):
                ei, gr = _ei_value_and_grad(X, model, plugin, type, True, proxy, device)   # R/max_crit.R:119
                return -ei, -np.asarray(gr)
            res = batched_lbfgsb(fg, starts, lower, upper, maxit=int(control.get("maxit", 100)))
            improved = -res["f"] >= vals
            starts = np.where(improved[:, None], res["x"], starts); vals = np.where(improved, -res["f"], vals)
        b = int(np.argmax(vals))
        return {"par": starts[b].copy(), "value": float(vals[b])}
    best = None
    for _ in range(restarts):
        x = lower + rng.random((N, d)) * (upper - lower)
        idx, val = screen_EI(x, model, plugin=None, type=type, minimization=minimization, proxy=proxy, device=device)
        par, value = x[idx[0]], val[0]
        if bfgs:
            # the reference polishes with minimization = TRUE whatever the caller asked (R/max_crit.R:119)
            res = minimize(lambda t: -EI(t, model, plugin=plugin, type=type, minimization=True, proxy=proxy, device=device)[0],
                           par, jac=lambda t: -EI_grad(t, model, plugin=plugin, type=type, minimization=True, proxy=proxy, device=device),
                           method="L-BFGS-B", bounds=list(zip(lower, upper)), options={"maxiter": 100})
            par, value = res.x, -res.fun
        if best is None or value > best[1]:
            best = (np.asarray(par, float), float(value))
    return {"par": best[0], "value": best[1]}


def max_EI(model: KM, plugin=None, type="UK", lower=None, upper=None, parinit=None, minimization=True, control=None,
           seed=None, device=0):
    """R/max_EI.R:185-219 with the GPU engine in place of rgenoud: the reference evolves a population of
    ``pop.size`` points for ``max.generations`` generations and hands the best to BFGS; here ``control["n.screen"]``
    uniform candidates (default 2^17, plus ``parinit``) are scored in one launch, the ``pop.size`` best are polished
    together by the batched L-BFGS (``control["BFGSmaxit"]`` iterations at most) and the best polished point is
    returned as list(par, value).  Same arguments and defaults as the reference (R/max_EI.R:195-204)."""
    from .optim import batched_lbfgsb
    control = dict(control or {})
    d = model.d
    lower = np.zeros(d) if lower is None else np.asarray(lower, float)
    upper = np.ones(d) if upper is None else np.asarray(upper, float)
    if plugin is None:
        plugin = float(np.min(model.y))                                   # R/max_EI.R:187 (not sign-flipped there)
    Npop = 3 * 2 ** d if d <= 6 else 32 * d
    pop = int(control.get("pop.size", Npop))
    maxit = int(control.get("BFGSmaxit", Npop))
    n_screen = int(control.get("n.screen", 1 << 17))
    rng = np.random.default_rng(seed)
    x = lower + rng.random((n_screen, d)) * (upper - lower)
    if parinit is not None:
        x[0] = np.clip(np.asarray(parinit, float), lower, upper)
    k = min(pop, n_screen)
    if k <= 64:                                                           # the device top-k keeps at most 64
        idx, val = screen_EI(x, model, plugin=plugin, type=type, minimization=minimization, k=k, device=device)
        idx = idx[idx >= 0]
    else:
        ei = EI(x, model, plugin=plugin, type=type, minimization=minimization, device=device)
        ei = np.where(np.isnan(ei), -np.inf, ei)
        idx = np.argpartition(-ei, k - 1)[:k]
        idx = idx[np.lexsort((idx, -ei[idx]))]                            # by value, first index wins ties
        val = ei[idx]
    starts = x[idx]

    if control.get("optimizer", "device") == "device":
        from .optim import device_polish
        res = device_polish(model, "EI", starts, lower, upper, type=type, minimization=minimization, plugin=plugin,
                            maxit=maxit, device=device)
        res = {"x": res["x"], "f": -res["f"]}
    else:
        def fg(X):
            ei, gr = _ei_value_and_grad(X, model, plugin, type, minimization, False, device)
            return -ei, -np.asarray(gr)
        res = batched_lbfgsb(fg, starts, lower, upper, maxit=maxit)
    vals = np.where(np.isfinite(res["f"]), -res["f"], -np.inf)
    b = int(np.argmax(vals))
    if not vals[b] >= val[0]:
        return {"par": x[idx[0]].copy(), "value": float(val[0])}
    return {"par": res["x"][b].copy(), "value": float(vals[b])}


def qEI(x, model: KM, plugin=None, type="UK", minimization=True, fastCompute=True, eps=1e-5, envir=None, device=0):
    """Multipoint expected improvement, R/qEI.R:119-194.  ``x`` is one batch (q x d) -> a scalar, or a
    stack of batches (B x q x d) -> B values.  The engine evaluates the exact second term for both
    values of ``fastCompute`` (the reference's fast path is a finite difference of it); ``eps`` is unused."""
    x = np.asarray(x, dtype=np.float64)
    single = x.ndim <= 2
    if x.ndim <= 1:
        x = x.reshape(-1, model.d)          # R: matrix(x, ncol = d)
    if x.ndim == 2:
        x = x[None]
    B, q, d = x.shape
    if d != model.d:
        raise ValueError("x does not have the right size")
    if plugin is None:                       # R/qEI.R:129-130 (max(y), not -max(y), when maximising)
        plugin = float(np.min(model.y)) if minimization else float(np.max(model.y))
    rows = np.asfortranarray(x.reshape(B * q, d))
    out = np.empty(B)
    h = model.handle(device)
    if envir is not None and minimization and 2 <= q:
        # qEI(x, envir) ... qEI.grad(x, envir) (R/qEI.R:190-198, R/qEI.grad.R:299-316): the reference stores the kriging moments
        # and CDF terms for the gradient call that optim issues at the same x.  Here one pass over the model gives value and
        # gradient (dk_qei_value_grad_dev, the value at x itself rather than at round(x, 8)); the gradient waits in envir.
        import torch
        with torch.cuda.device(h.device):
            xd = torch.from_numpy(np.ascontiguousarray(rows.T)).cuda()
            vd = torch.empty(B, dtype=torch.float64, device="cuda"); gd = torch.empty((d, B * q), dtype=torch.float64, device="cuda")
            _lib.check(_lib.load().dk_qei_value_grad_dev(h.ptr, C.c_void_p(xd.data_ptr()), B, q, float(plugin), _type_id(type), 1,
                                                         C.c_void_p(vd.data_ptr()), C.c_void_p(gd.data_ptr()), None),
                       "dk_qei_value_grad_dev")
            torch.cuda.synchronize()
            out = vd.cpu().numpy()
            envir.update({"qEI.x": rows.copy(), "qEI.args": (float(plugin), type, q),
                          "qEI.grad": np.ascontiguousarray(gd.cpu().numpy().T).reshape(B, q, d)})
        return float(out[0]) if single else out
    _lib.check(_lib.load().dk_qei(h.ptr, _vp(rows), B, q, float(plugin), _type_id(type), int(bool(minimization)),
                                  _vp(out)), "dk_qei")
    return float(out[0]) if single else out


def AKG(x, model: KM, new_noise_var=0.0, type="UK", envir=None, device=0):
    """Approximate knowledge gradient, R/AKG.R:93-229.  A length-d ``x`` gives a scalar (the reference's
    contract); an N x d matrix gives N values from one launch (the noisy DOE scans of
    R/noisy.optimizer.R:456-457 become one call)."""
    single = np.asarray(x).ndim <= 1
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    out = np.empty(N)
    h = model.handle(device)
    if envir is not None:
        # AKG(x, envir) ... AKG.grad(x, envir) (R/AKG.R:221-228 stores its intermediate terms for R/AKG.grad.R:84-94): the
        # gradient kernel returns the value too, so one launch serves both calls; the gradient waits in envir
        grad = np.empty((N, model.d), order="F")
        _lib.check(_lib.load().dk_akg_grad(h.ptr, _vp(x), N, float(new_noise_var), _type_id(type), _vp(grad), _vp(out)), "dk_akg_grad")
        envir.update({"AKG.x": np.array(x, copy=True), "AKG.args": (float(new_noise_var), type), "AKG.grad": grad})
        return float(out[0]) if single else out
    _lib.check(_lib.load().dk_akg(h.ptr, _vp(x), N, float(new_noise_var), _type_id(type), _vp(out)), "dk_akg")
    return float(out[0]) if single else out


def randLHS(n, k, rng):
    """R/sampleFromEI.R:143-149: random Latin hypercube (numpy RNG instead of R's runif)."""
    perm = np.argsort(rng.random((n, k)), axis=0)
    return (perm + rng.random((n, k))) / n


def sampleFromEI(model: KM, minimization=True, n=1, initdistrib=None, lower=None, upper=None, T=None, seed=None,
                 device=0, method=None):
    """R/sampleFromEI.R:77-141: draw ``n`` points with probability proportional to EI among 1000 d LHS points
    (or ``initdistrib``).  method = "device" (default for n <= 64): weights, the reference's fix-up and the draw without
    replacement all on the GPU (dk_sample_from_ei, counter-based uniforms seeded by ``seed``); "host": weights on the GPU,
    numpy's generator for the draw in place of R's sample()."""
    d = model.d
    lower = np.zeros(d) if lower is None else np.asarray(lower, float)
    upper = np.ones(d) if upper is None else np.asarray(upper, float)
    if lower.size != upper.size:
        raise ValueError("Error in samplefromEI : lower and upper must have the same length")
    rng = np.random.default_rng(seed)
    pts = lower + randLHS(1000 * d, d, rng) * (upper - lower) if initdistrib is None else np.asarray(initdistrib, float)
    pts = _as_candidates(pts, d)
    N = pts.shape[0]
    if T is None:
        T = float(np.min(model.y)) if minimization else float(np.max(model.y))
    h = model.handle(device)
    if method is None:
        method = "device" if n <= 64 else "host"
    if method == "device":
        idx = np.empty(n, dtype=np.int64)
        s64 = int(rng.integers(0, 2 ** 63)) if seed is None else int(seed) & (2 ** 64 - 1)
        _lib.check(_lib.load().dk_sample_from_ei(h.ptr, _vp(pts), N, float(T), int(bool(minimization)), int(n), s64, _vp(idx)),
                   "dk_sample_from_ei")
        return np.asarray(pts[idx]).reshape(n, d)
    ei = np.empty(N)
    _lib.check(_lib.load().dk_sample_weights(h.ptr, _vp(pts), N, float(T), int(bool(minimization)), _vp(ei)),
               "dk_sample_weights")
    if np.count_nonzero(ei) < n:                             # R/sampleFromEI.R:126-133
        mx = ei.max()
        ei = ei / mx + 1.0 / (1000 * d) if mx > 0 else np.ones(N)
    idx = rng.choice(N, size=n, replace=False, p=ei / ei.sum())
    return np.asarray(pts[idx]).reshape(n, d)


def qEI_grad(x, model: KM, plugin=None, type="UK", minimization=True, fastCompute=True, eps=1e-6, envir=None, device=0):
    """Gradient of qEI with respect to the batch, R/qEI.grad.R:267-424: q x d for one batch, B x q x d for a
    stack.  q = 1 is routed to EI.grad as the reference does."""
    x = np.asarray(x, dtype=np.float64)
    single = x.ndim <= 2
    if x.ndim <= 1:
        x = x.reshape(-1, model.d)
    if x.ndim == 2:
        x = x[None]
    B, q, d = x.shape
    if d != model.d:
        raise ValueError("x does not have the right size")
    if plugin is None:                       # R/qEI.grad.R:272-276, 297
        plugin = float(np.min(model.y)) if minimization else -float(np.max(model.y))
    if q == 1:
        if not minimization:
            raise ValueError("qEI.grad doesn't work in 'minimization = FALSE' when dim = 1 (in progress).")
        g = np.asarray(EI_grad(x[:, 0, :], model, plugin=plugin, type=type, device=device)).reshape(B, 1, d)
        return g[0] if single else g
    rows = np.asfortranarray(x.reshape(B * q, d))
    if (envir is not None and minimization and "qEI.grad" in envir and envir.get("qEI.args") == (float(plugin), type, q)
            and np.array_equal(envir["qEI.x"], rows)):
        g = envir["qEI.grad"]                # left there by qEI(x, envir) at this very x
        return g[0] if single else g
    out = np.empty((B * q, d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_qei_grad(h.ptr, _vp(rows), B, q, float(plugin), _type_id(type), int(bool(minimization)),
                                       _vp(out)), "dk_qei_grad")
    g = np.ascontiguousarray(out).reshape(B, q, d)
    return g[0] if single else g


def max_qEI(model: KM, npoints, lower, upper, crit="exact", minimization=True, optimcontrol=None, seed=None, device=0):
    """R/max_qEI.R:176-225, crit = "exact", optimcontrol$method = "BFGS": nStarts L-BFGS-B runs of qEI / qEI.grad
    from batches drawn by sampleFromEI; the best run wins.  All nStarts start batches are scored in ONE qEI launch
    first.  crit = "CL" (constant liar) and method = "genoud" need DiceKriging's model update / rgenoud and stay
    outside the engine."""
    from scipy.optimize import minimize
    if crit == "CL":
        return _max_qEI_CL(model, npoints, (optimcontrol or {}).get("L"), lower, upper,
                           parinit=(optimcontrol or {}).get("parinit"), minimization=minimization,
                           control=optimcontrol, seed=seed, device=device)
    if crit != "exact":
        raise ValueError(f"'{crit}' is unknown.")
    oc = dict(optimcontrol or {})
    if oc.get("method", "BFGS") != "BFGS":
        raise NotImplementedError("only optimcontrol$method = 'BFGS' is served")
    d = model.d
    lower = np.asarray(lower, float); upper = np.asarray(upper, float)
    n_starts = int(oc.get("nStarts", 4))
    fast = bool(oc.get("fastCompute", True))
    start = oc.get("parinit")
    if start is None:                                         # R/max_qEI.R:204
        start = sampleFromEI(model, minimization=minimization, n=npoints * n_starts, lower=lower, upper=upper,
                             seed=seed, device=device).reshape(n_starts, npoints, d)
    start = np.asarray(start, float).reshape(-1, npoints, d)
    lo = np.tile(lower, npoints); up = np.tile(upper, npoints)
    plugin_v = float(np.min(model.y)) if minimization else float(np.max(model.y))
    plugin_g = float(np.min(model.y)) if minimization else -float(np.max(model.y))
    optimizer = oc.get("optimizer", "device" if (minimization and 2 <= npoints <= 10) else "batched")
    if optimizer == "device":
        # all nStarts runs advance together ON THE DEVICE (dk_polish): one qEI + one qEI.grad launch per trip.  (Maximisation
        # keeps the host variant: the reference evaluates value and gradient with different plugins there, R/qEI.R:130 vs
        # R/qEI.grad.R:272-276.)
        from .optim import device_polish
        S = start.shape[0]
        res = device_polish(model, "qEI", start.reshape(S, npoints * d), lower, upper, q=npoints, minimization=True,
                            plugin=plugin_v, maxit=int(oc.get("maxit", 100)), device=device)
        vals = np.where(np.isfinite(res["f"]), res["f"], -np.inf)
        b = int(np.argmax(vals))
        return {"par": res["x"][b].reshape(npoints, d).copy(), "value": float(vals[b])}
    if optimizer == "batched":
        # all nStarts runs advance together: one qEI and one qEI.grad launch per evaluation (optim.py)
        from .optim import batched_lbfgsb
        S = start.shape[0]

        def fg(X):
            xb = X.reshape(S, npoints, d)
            val = np.atleast_1d(qEI(xb, model, plugin=plugin_v, minimization=minimization, fastCompute=fast, device=device))
            gr = qEI_grad(xb, model, plugin=plugin_g, minimization=minimization, fastCompute=fast, device=device)
            return -val, -np.asarray(gr).reshape(S, npoints * d)
        res = batched_lbfgsb(fg, start.reshape(S, npoints * d), lo, up, maxit=int(oc.get("maxit", 100)))
        vals = np.where(np.isfinite(res["f"]), -res["f"], -np.inf)
        b = int(np.argmax(vals))
        return {"par": res["x"][b].reshape(npoints, d).copy(), "value": float(vals[b])}
    best = None
    for s in range(start.shape[0]):
        f = lambda t: -qEI(t.reshape(npoints, d), model, plugin=plugin_v, minimization=minimization, fastCompute=fast, device=device)
        g = lambda t: -qEI_grad(t.reshape(npoints, d), model, plugin=plugin_g, minimization=minimization, fastCompute=fast,
                                device=device).ravel()
        res = minimize(f, start[s].ravel(), jac=g, method="L-BFGS-B", bounds=list(zip(lo, up)),
                       options={"maxiter": int(oc.get("maxit", 100))})
        if best is None or -res.fun > best[1]:
            best = (res.x.reshape(npoints, d), float(-res.fun))
    return {"par": best[0], "value": best[1]}


def AKG_grad(x, model: KM, new_noise_var=0.0, type="UK", envir=None, device=0):
    """Gradient of AKG, R/AKG.grad.R:65-281: length d for one point, N x d for a matrix.  With the ``envir`` that
    ``AKG(x, envir)`` filled at the same x the stored gradient is returned (the reference reuses its stored terms)."""
    single = np.asarray(x).ndim <= 1
    x = _as_candidates(x, model.d)
    N = x.shape[0]
    if (envir is not None and "AKG.grad" in envir and envir.get("AKG.args") == (float(new_noise_var), type)
            and np.array_equal(envir["AKG.x"], x)):
        grad = envir["AKG.grad"]             # left there by AKG(x, envir) at this very x
        return grad[0].copy() if single else grad
    grad = np.empty((N, model.d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_akg_grad(h.ptr, _vp(x), N, float(new_noise_var), _type_id(type), _vp(grad), None), "dk_akg_grad")
    return grad[0].copy() if single else grad


def AEI(x, model: KM, new_noise_var=0.0, y_min=None, type="UK", envir=None, device=0):
    """Augmented expected improvement, R/AEI.R:105-146 (rows batched)."""
    single = np.asarray(x).ndim <= 1
    if y_min is None:                                        # R/AEI.R:116-122
        pr = predict(model, model.X, type=type, device=device)
        y_min = float(pr["mean"][int(np.argmin(pr["mean"] + ndtri(0.75) * pr["sd"]))])
    x = _as_candidates(x, model.d)
    out = np.empty(x.shape[0])
    h = model.handle(device)
    _lib.check(_lib.load().dk_aei(h.ptr, _vp(x), x.shape[0], float(y_min), float(new_noise_var), _type_id(type), _vp(out)), "dk_aei")
    return float(out[0]) if single else out


def EQI(x, model: KM, new_noise_var=0.0, beta=0.9, q_min=None, type="UK", envir=None, device=0):
    """Expected quantile improvement, R/EQI.R:100-151 (rows batched)."""
    single = np.asarray(x).ndim <= 1
    if q_min is None:                                        # R/EQI.R:110-112
        pr = predict(model, model.X, type=type, device=device)
        q_min = float(np.min(pr["mean"] + ndtri(beta) * pr["sd"]))
    x = _as_candidates(x, model.d)
    out = np.empty(x.shape[0])
    h = model.handle(device)
    _lib.check(_lib.load().dk_eqi(h.ptr, _vp(x), x.shape[0], float(q_min), float(new_noise_var), float(ndtri(beta)),
                                  _type_id(type), _vp(out)), "dk_eqi")
    return float(out[0]) if single else out


def _default_y_min(model, type, device):
    pr = predict(model, model.X, type=type, device=device)          # R/AEI.R:116-122
    return float(pr["mean"][int(np.argmin(pr["mean"] + ndtri(0.75) * pr["sd"]))])


def AEI_grad(x, model: KM, new_noise_var=0.0, y_min=None, type="UK", envir=None, device=0):
    """Gradient of AEI, R/AEI.grad.R:65-154 (rows batched)."""
    single = np.asarray(x).ndim <= 1
    if y_min is None:
        y_min = _default_y_min(model, type, device)
    x = _as_candidates(x, model.d)
    g = np.empty((x.shape[0], model.d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_aei_grad(h.ptr, _vp(x), x.shape[0], float(y_min), float(new_noise_var), _type_id(type), _vp(g)),
               "dk_aei_grad")
    return g[0].copy() if single else g


def EQI_grad(x, model: KM, new_noise_var=0.0, beta=0.9, q_min=None, type="UK", envir=None, device=0):
    """Gradient of EQI, R/EQI.grad.R:66-146 (rows batched)."""
    single = np.asarray(x).ndim <= 1
    if q_min is None:
        pr = predict(model, model.X, type=type, device=device)
        q_min = float(np.min(pr["mean"] + ndtri(beta) * pr["sd"]))
    x = _as_candidates(x, model.d)
    g = np.empty((x.shape[0], model.d), order="F")
    h = model.handle(device)
    _lib.check(_lib.load().dk_eqi_grad(h.ptr, _vp(x), x.shape[0], float(q_min), float(new_noise_var), float(ndtri(beta)),
                                       _type_id(type), _vp(g)), "dk_eqi_grad")
    return g[0].copy() if single else g


# ---- constrained criteria over several km models (objective + constraints), SURVEY 8(f) rank 3 ----
_DISTANCES = {"euclidean": 0, "covdist": 1, "covratio": 2}


def _handles(models, device):
    hs = [m.handle(device) for m in models]
    arr = (C.c_void_p * len(hs))(*[h.ptr for h in hs])
    return hs, arr


def _model_list(model):
    if isinstance(model, KM):
        return [model]
    out = list(model)
    if not out or not all(isinstance(m, KM) for m in out):
        raise TypeError("model must be a km model or a list of km models (fastfun objects are not served by the GPU engine)")
    return out


def test_feas_vec(cst, equality=False, tolConstraints=None):
    """R/test_feas_vec.R:11-36: which rows of the observed constraint matrix are feasible."""
    cst = np.asarray(cst, float)
    if cst.ndim == 1:
        cst = cst.reshape(1, -1)
    nc = cst.shape[1]
    eq = np.broadcast_to(np.asarray(equality, bool), (nc,)) if np.ndim(equality) else np.full(nc, bool(equality))
    if tolConstraints is None:
        eps = np.zeros(nc); eps[eq] = 0.05
    else:
        eps = np.broadcast_to(np.asarray(tolConstraints, float), (nc,))
    ok = np.ones(cst.shape[0], bool)
    if eq.any():
        ok[np.max(cst[:, eq] + eps[eq], axis=1) < 0] = False
    ok[np.max(cst - eps, axis=1) > 0] = False
    return ok


test_feas_vec.__test__ = False   # a reference function name, not a pytest case


def checkPredict(x, model, threshold=1e-4, distance="covdist", type="UK", device=0, return_distance=False):
    """R/checkPredict.R:18-75 for every row of ``x``: TRUE where the point is too close to the design of one of the
    models.  ``return_distance=True`` gives the distances themselves (the reference's ``mindist``)."""
    models = _model_list(model)
    if distance is None:
        distance = "covdist"
    if threshold is None:
        threshold = 1e-4
    x = _as_candidates(x, models[0].d)
    N = x.shape[0]
    if distance == "none":
        return np.zeros(N, bool)
    if distance not in _DISTANCES:
        raise ValueError(f"distance must be one of euclidean, covdist, covratio, none; got {distance!r}")
    out = np.empty(N)
    hs, arr = _handles(models, device)
    _lib.check(_lib.load().dk_check_predict(arr, len(hs), _vp(x), N, _type_id(type), _DISTANCES[distance], _vp(out)),
               "dk_check_predict")
    return out if return_distance else out < threshold


def Feas(model: KM, newdata, type="UK", equality=False, threshold=0.0, device=0):
    """Probability of feasibility, R/crit_EFI.R:231-238 (the reference predicts with "UK" whatever ``type`` says)."""
    res = predict(model, newdata, type="UK", device=device)
    with np.errstate(divide="ignore", invalid="ignore"):
        hi = np.where(res["sd"] > 0, ndtr((threshold - res["mean"]) / res["sd"]), (threshold >= res["mean"]) * 1.0)
        if not equality:
            return hi
        lo = np.where(res["sd"] > 0, ndtr((-threshold - res["mean"]) / res["sd"]), (-threshold >= res["mean"]) * 1.0)
    return hi - lo


def _constraint_setup(model_fun, model_constraint, equality, critcontrol):
    cst = _model_list(model_constraint)
    nc = len(cst)
    cc = dict(critcontrol or {})
    eqv = np.full(nc, False) if (np.ndim(equality) == 0 and not equality) else np.broadcast_to(np.asarray(equality, bool), (nc,)).copy()
    observations = np.column_stack([m.y for m in cst])
    feas = test_feas_vec(observations, equality=eqv if eqv.any() else False, tolConstraints=cc.get("tolConstraints"))
    best = float(np.min(model_fun.y[feas])) if feas.any() else np.inf
    if cc.get("tolConstraints") is None:
        tol = np.zeros(nc); tol[eqv] = 0.05
    else:
        tol = np.broadcast_to(np.asarray(cc["tolConstraints"], float), (nc,)).copy()
    return cst, cc, eqv, tol, best


def _feasible_improvement(x, model_fun, cst, plugin, type, kind, eqv, tol, device):
    x = _as_candidates(x, model_fun.d)
    N = x.shape[0]
    out = np.empty(N)
    hf = model_fun.handle(device)
    hs, arr = _handles(cst, device)
    eq = np.ascontiguousarray(eqv, dtype=np.int32)
    tl = np.ascontiguousarray(tol, dtype=np.float64)
    _lib.check(_lib.load().dk_feasible_improvement(hf.ptr, arr, len(hs), _vp(x), N, float(plugin), _type_id(type), kind,
                                                   _vp(eq), _vp(tl), _vp(out)), "dk_feasible_improvement")
    return out


def crit_EFI(x, model_fun: KM, model_constraint, equality=False, critcontrol=None, plugin=None, type="UK", device=0):
    """Expected feasible improvement, R/crit_EFI.R:125-228, for every row of ``x``: 0 where checkPredict refuses the
    point, else EI(x; model.fun, plugin = best feasible observation) x prod_i Feas_i(x)."""
    if not isinstance(model_fun, KM):
        raise TypeError("model.fun must be a km model (fastfun objectives are not served by the GPU engine)")
    cst, cc, eqv, tol, best = _constraint_setup(model_fun, model_constraint, equality, critcontrol)
    if plugin is None:
        plugin = best
    single = np.asarray(x).ndim <= 1
    out = _feasible_improvement(x, model_fun, cst, plugin, type, 0, eqv, tol, device)
    refuse = checkPredict(x, cst + [model_fun], threshold=cc.get("threshold"), distance=cc.get("distance"), type=type,
                          device=device)
    out[refuse] = 0.0
    return float(out[0]) if single else out


def prob_of_feasible_improvement(integration_points, model_fun: KM, model_constraint, equality=False, critcontrol=None,
                                 type="UK", device=0):
    """R/prob.of.feasible.improvement.R:1-81: pnorm((obj.min - mean) / sd) x prod_i Feas_i over the integration
    points; inequality thresholds are 0, equality tolerances default to 0.05."""
    cst, cc, eqv, tol, best = _constraint_setup(model_fun, model_constraint, equality, critcontrol)
    thr = np.where(eqv, tol, 0.0)
    return _feasible_improvement(integration_points, model_fun, cst, best, type, 1, eqv, thr, device)


def _max_qEI_CL(model: KM, npoints, L, lower, upper, parinit=None, minimization=True, control=None, seed=None, device=0):
    """max_qEI.CL, R/max_qEI.R:283-318: npoints rounds of max_EI, after each of which the maximiser joins the design
    with a "lie" as response (L = NULL: min(y) or max(y); "min"; "max"; "mean" = kriging believer; or a number) and the
    model's T, z, M are refreshed with covariance and trend coefficients kept.  The refresh is the n^2 append of
    ``update_km`` instead of computeAuxVariables' full Cholesky, and the device handle grows IN PLACE (dk_model_append on a
    device-to-device replica of the caller's handle: per round only the last block rows of the tiled factor are rebuilt)."""
    from .update import update_km
    n1 = model.n
    KB = False
    if L is None:
        liar = float(np.min(model.y)) if minimization else float(np.max(model.y))
    elif isinstance(L, str):
        if L == "max":
            liar = float(np.max(model.y))
        elif L == "min":
            liar = float(np.min(model.y))
        elif L == "mean":
            KB = True
        else:
            raise ValueError('control$L must be NULL, "max", "min", "mean" or a scalar specifying the plugin.')
    else:
        liar = float(L)
    ctrl = {k: v for k, v in dict(control or {}).items() if k in ("pop.size", "BFGSmaxit", "n.screen")}
    parinit = None if parinit is None else np.atleast_2d(np.asarray(parinit, float))     # d x npoints as in R
    rng = np.random.default_rng(seed)
    cur = model
    created = []
    for s in range(npoints):
        start = parinit[:, s] if (parinit is not None and s < parinit.shape[1]) else None
        o = max_EI(cur, lower=lower, upper=upper, parinit=start, minimization=minimization, control=ctrl,
                   seed=int(rng.integers(1 << 31)), device=device)
        if KB:
            liar = float(predict(cur, o["par"], device=device)["mean"][0])
        nxt = update_km(cur, o["par"], liar, new_noise_var=0.0, reuse_handles="clone" if cur is model else True)
        if cur is not model:
            cur.release()
        cur = nxt
        created.append(nxt)
    par, value = cur.X[n1:n1 + npoints].copy(), cur.y[n1:n1 + npoints].reshape(-1, 1).copy()
    if cur is not model:
        cur.release()
    return {"par": par, "value": value}


def crit_SUR_cst(x, model_fun: KM, model_constraint, equality=False, critcontrol=None, type="UK", device=0):
    """Stepwise uncertainty reduction with constraints, R/crit_SUR_cst.R:120-378, for every row of ``x``.
    ``critcontrol["integration.points"]`` (M x d) or, to have them drawn by integration_design_cst, ``lower`` and
    ``upper``; ``integration.weights`` defaults to 1 / M each.  0 where checkPredict refuses the point; NaN when there are
    more than three constraints (the reference warns and returns NA)."""
    if not isinstance(model_fun, KM):
        raise TypeError("model.fun must be a km model (fastfun objectives are not served by the GPU engine)")
    cst, cc, eqv, tol, best = _constraint_setup(model_fun, model_constraint, equality, critcontrol)
    single = np.asarray(x).ndim <= 1
    x = _as_candidates(x, model_fun.d)
    N = x.shape[0]
    if len(cst) > 3:
        import warnings
        warnings.warn("crit_SUR_cst does not take more than 3 constraints")
        return float("nan") if single else np.full(N, np.nan)
    if cc.get("integration.points") is None:
        # R/crit_SUR_cst.R:139-143: integration_design_cst(integcontrol = critcontrol, lower, upper, ...)
        if cc.get("lower") is None or cc.get("upper") is None:
            raise ValueError("critcontrol needs 'integration.points', or 'lower' and 'upper' to draw them")
        des = integration_design_cst(integcontrol=cc, lower=cc["lower"], upper=cc["upper"], model_fun=model_fun,
                                     model_constraint=cst, device=device)
        cc["integration.points"] = des["integration.points"]
        cc["integration.weights"] = des["integration.weights"]
    P = np.asfortranarray(np.atleast_2d(np.asarray(cc["integration.points"], float)))
    if P.shape[1] != model_fun.d:
        raise ValueError("integration.points does not have the right size")
    w = cc.get("integration.weights")
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    if w is not None and w.shape != (P.shape[0],):
        raise ValueError("integration.weights must have one entry per integration point")
    out = np.empty(N)
    hf = model_fun.handle(device)
    hs, arr = _handles(cst, device)
    eq = np.ascontiguousarray(eqv, dtype=np.int32)
    tl = np.ascontiguousarray(tol, dtype=np.float64)
    _lib.check(_lib.load().dk_crit_sur_cst(hf.ptr, arr, len(hs), _vp(x), N, _vp(P), P.shape[0], _vp(w), float(best),
                                           _type_id(type), _vp(eq), _vp(tl), _vp(out)), "dk_crit_sur_cst")
    refuse = checkPredict(x, cst + [model_fun], threshold=cc.get("threshold"), distance=cc.get("distance"), type=type,
                          device=device)
    out[refuse] = 0.0
    return float(out[0]) if single else out


# ---- the noisy-criteria maximisers (rgenoud in the reference): screen, then batched polish of the best starts ----
def _screen_and_polish(value, value_grad, d, lower, upper, parinit, control, seed, maximise=True, device_crit=None):
    """Engine replacement of the genoud calls of R/max_AEI.R:95-112, R/max_AKG.R:97-113, R/max_EQI.R:97-114 and
    R/min_quantile.R:96-108: ``control["n.screen"]`` uniform candidates (default 2^15, plus ``parinit``) are scored in
    one launch, the ``pop.size`` best (reference default 3 x 2^d for d <= 6, else 32 d) are polished together by the
    batched projected L-BFGS, the best polished point wins.  Returns list(par, value) as the reference does."""
    from .optim import batched_lbfgsb
    control = dict(control or {})
    lower = np.zeros(d) if lower is None else np.asarray(lower, float)
    upper = np.ones(d) if upper is None else np.asarray(upper, float)
    Npop = 3 * 2 ** d if d <= 6 else 32 * d
    pop = int(control.get("pop.size", Npop))
    maxit = int(control.get("BFGSmaxit", Npop))
    n_screen = int(control.get("n.screen", 1 << 15))
    rng = np.random.default_rng(seed)
    x = lower + rng.random((n_screen, d)) * (upper - lower)
    if parinit is not None:
        x[0] = np.clip(np.asarray(parinit, float), lower, upper)
    sgn = -1.0 if maximise else 1.0
    v = sgn * np.asarray(value(x), float)
    v = np.where(np.isnan(v), np.inf, v)
    k = min(pop, n_screen)
    idx = np.argpartition(v, k - 1)[:k]
    idx = idx[np.lexsort((idx, v[idx]))]

    if device_crit is not None and control.get("optimizer", "device") == "device":
        # the polish on the device: device_crit = (model, criterion name, keyword arguments of optim.device_polish)
        from .optim import device_polish
        mdl, name, kw = device_crit
        res = device_polish(mdl, name, x[idx], lower, upper, maxit=maxit, **kw)
        res = {"x": res["x"], "f": sgn * res["f"]}
    else:
        def fg(X):
            f, g = value_grad(X)
            return sgn * np.asarray(f, float), sgn * np.asarray(g, float)
        res = batched_lbfgsb(fg, x[idx], lower, upper, maxit=maxit)
    f = np.where(np.isfinite(res["f"]), res["f"], np.inf)
    b = int(np.argmin(f))
    if not f[b] <= v[idx[0]]:
        return {"par": x[idx[0]].reshape(1, -1).copy(), "value": float(sgn * v[idx[0]])}
    return {"par": res["x"][b].reshape(1, -1).copy(), "value": float(sgn * f[b])}


def max_AEI(model: KM, new_noise_var=0.0, y_min=None, type="UK", lower=None, upper=None, parinit=None, control=None,
            seed=None, device=0):
    """R/max_AEI.R:79-118."""
    if y_min is None:
        y_min = _default_y_min(model, type, device)
    return _screen_and_polish(
        lambda X: AEI(X, model, new_noise_var=new_noise_var, y_min=y_min, type=type, device=device),
        lambda X: (AEI(X, model, new_noise_var=new_noise_var, y_min=y_min, type=type, device=device),
                   AEI_grad(X, model, new_noise_var=new_noise_var, y_min=y_min, type=type, device=device)),
        model.d, lower, upper, parinit, control, seed,
        device_crit=(model, "AEI", dict(type=type, plugin=y_min, new_noise_var=new_noise_var, device=device)))


def max_AKG(model: KM, new_noise_var=0.0, type="UK", lower=None, upper=None, parinit=None, control=None, seed=None,
            device=0):
    """R/max_AKG.R:82-120."""
    return _screen_and_polish(
        lambda X: AKG(X, model, new_noise_var=new_noise_var, type=type, device=device),
        lambda X: (AKG(X, model, new_noise_var=new_noise_var, type=type, device=device),
                   AKG_grad(X, model, new_noise_var=new_noise_var, type=type, device=device)),
        model.d, lower, upper, parinit, control, seed,
        device_crit=(model, "AKG", dict(type=type, new_noise_var=new_noise_var, device=device)))


def max_EQI(model: KM, new_noise_var=0.0, beta=0.9, q_min=None, type="UK", lower=None, upper=None, parinit=None,
            control=None, seed=None, device=0):
    """R/max_EQI.R:80-120."""
    if q_min is None:
        p = predict(model, model.X, type=type, device=device)
        q_min = float(np.min(p["mean"] + ndtri(beta) * p["sd"]))            # R/EQI.R:108-112
    return _screen_and_polish(
        lambda X: EQI(X, model, new_noise_var=new_noise_var, beta=beta, q_min=q_min, type=type, device=device),
        lambda X: (EQI(X, model, new_noise_var=new_noise_var, beta=beta, q_min=q_min, type=type, device=device),
                   EQI_grad(X, model, new_noise_var=new_noise_var, beta=beta, q_min=q_min, type=type, device=device)),
        model.d, lower, upper, parinit, control, seed,
        device_crit=(model, "EQI", dict(type=type, plugin=q_min, new_noise_var=new_noise_var, qnorm_beta=float(ndtri(beta)),
                                        device=device)))


def min_quantile(model: KM, beta=0.1, type="UK", lower=None, upper=None, parinit=None, control=None, seed=None, device=0):
    """R/min_quantile.R:77-115 (a minimisation)."""
    return _screen_and_polish(
        lambda X: kriging_quantile(X, model, beta=beta, type=type, device=device),
        lambda X: (kriging_quantile(X, model, beta=beta, type=type, device=device),
                   kriging_quantile_grad(X, model, beta=beta, type=type, device=device)),
        model.d, lower, upper, parinit, control, seed, maximise=False,
        device_crit=(model, "quantile", dict(type=type, qnorm_beta=float(ndtri(beta)), device=device)))


def integration_design_cst(integcontrol=None, lower=None, upper=None, model_fun=None, model_constraint=None, equality=False,
                           critcontrol=None, min_prob=0.001, seed=None, device=0):
    """R/integration_design_cst.R: the integration points (and weights) of crit_SUR_cst.  No ``integration.points`` in
    ``integcontrol`` -> 100 d uniform points; given -> passed through; otherwise ``distrib`` "MC" (uniform), "sobol"
    (scipy's Sobol sequence in place of randtoolbox) or "SUR" (importance sampling of ``n.candidates`` uniform points
    by prob.of.feasible.improvement, evaluated on the GPU).  Returns dict(integration.points, integration.weights)."""
    ic = dict(integcontrol or {})
    lower = np.asarray(lower, float); upper = np.asarray(upper, float)
    if lower.shape != upper.shape:
        raise ValueError("'lower' and 'upper' must have the same length")
    d = lower.size
    rng = np.random.default_rng(seed)
    if ic.get("integration.points") is not None:
        return {"integration.points": np.atleast_2d(np.asarray(ic["integration.points"], float)),
                "integration.weights": ic.get("integration.weights")}
    if not ic or ("n.points" not in ic and "distrib" not in ic):
        return {"integration.points": lower + rng.random((100 * d, d)) * (upper - lower), "integration.weights": None}
    n_points = int(ic.get("n.points", 100 * d))
    distrib = ic.get("distrib", "MC")

    def draw(kind, n):
        if kind == "sobol":
            from scipy.stats import qmc
            return lower + qmc.Sobol(d, scramble=False).random(n) * (upper - lower)
        if kind == "MC":
            return lower + rng.random((n, d)) * (upper - lower)
        raise ValueError(f"unknown distribution {kind!r}")
    if distrib in ("sobol", "MC"):
        return {"integration.points": draw(distrib, n_points), "integration.weights": None}
    if distrib != "SUR":
        raise ValueError(f"unknown integcontrol$distrib {distrib!r}")
    n_cand = int(ic.get("n.candidates", 10 * n_points))
    init = ic.get("init.distrib", "MC")
    cand = np.atleast_2d(np.asarray(ic["init.distrib.spec"], float)) if init == "spec" else draw(init, n_cand)
    if model_fun is None or model_constraint is None:
        raise ValueError("for the 'SUR' importance sampling distribution model.fun and model.constraint are needed")
    tau = prob_of_feasible_improvement(cand, model_fun, model_constraint, equality=equality, critcontrol=critcontrol,
                                       type="UK", device=device)
    tot = float(np.sum(tau)) or 1.0
    prob = np.maximum(tau / tot, min_prob / n_cand)
    prob = prob / prob.sum()
    weight = 1.0 / (prob * n_cand * n_points)
    cum = np.concatenate([[0.0], np.cumsum(prob)])
    idx = np.clip(np.searchsorted(cum, rng.random(n_points), side="right") - 1, 0, n_cand - 1)   # findInterval(all.inside)
    return {"integration.points": cand[idx], "integration.weights": weight[idx]}
