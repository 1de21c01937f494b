"""Batched box-constrained quasi-Newton minimiser for the multistart polish steps of the reference.

The reference polishes every restart with ``optim(method = "L-BFGS-B")`` one after the other, one criterion call per
function evaluation (R/max_crit.R:114-122, R/max_qEI.R:204-222).  On the GPU a criterion call costs the same for one
point as for a few thousand, so all restarts advance together here: every iteration is ONE batched value + gradient
launch for all S starts (plus one per backtracking step), and the quasi-Newton algebra on the S x D iterates is a few
vectorised numpy lines on the host.

Method per start: projected L-BFGS.  The variables at a bound whose gradient points outward are frozen, the two-loop
recursion runs on the free ones, the step is projected back into the box and shortened by Armijo backtracking along
the projection arc; a direction that is not a descent direction falls back to the projected gradient.  Stops per
start on the projected-gradient norm (``pgtol``), on the relative decrease (``factr`` x machine epsilon, as
``optim``'s L-BFGS-B does) or at ``maxit``.  The iterates differ from Nocedal's L-BFGS-B (no generalised Cauchy
point), the stationarity conditions reached are the same.
"""
from __future__ import annotations

import numpy as np

_EPS = np.finfo(float).eps


def batched_lbfgsb(fg, x0, lower, upper, maxit=100, m=8, pgtol=0.0, factr=1e7, c1=1e-4, max_backtracks=25):
    """Minimise S independent problems min f_s(x), lower <= x <= upper.

    fg(X) takes an S x D array and returns (f[S], g[S, D]) for ALL rows (rows that have converged are still passed:
    a fixed batch shape keeps one launch per call).  Returns dict(x, f, iterations, evaluations, converged)."""
    x = np.clip(np.asarray(x0, float), lower, upper)
    S, D = x.shape
    lower = np.broadcast_to(np.asarray(lower, float), (D,))
    upper = np.broadcast_to(np.asarray(upper, float), (D,))
    f, g = fg(x)
    f = np.array(f, float); g = np.array(g, float)
    bad = ~np.isfinite(f) | ~np.all(np.isfinite(g), axis=1)
    g[bad] = 0.0
    evals = 1
    sh = np.zeros((m, S, D)); yh = np.zeros((m, S, D)); rho = np.zeros((m, S))
    done = bad.copy()
    iters = np.zeros(S, dtype=int)
    for it in range(maxit):
        at_lo = (x <= lower) & (g > 0.0)
        at_up = (x >= upper) & (g < 0.0)
        free = ~(at_lo | at_up)
        pg = np.where(free, g, 0.0)
        done |= np.max(np.abs(pg), axis=1) <= pgtol
        if done.all():
            break
        # two-loop recursion on the free variables; pairs with rho = 0 are inert
        q = pg.copy()
        alpha = np.zeros((m, S))
        order = [(it - 1 - j) % m for j in range(min(it, m))]          # newest first
        for j in order:
            alpha[j] = rho[j] * np.einsum("sd,sd->s", sh[j], q)
            q -= alpha[j][:, None] * yh[j]
        if order:
            j0 = order[0]
            yy = np.einsum("sd,sd->s", yh[j0], yh[j0])
            gamma = np.where((rho[j0] > 0.0) & (yy > 0.0), 1.0 / np.maximum(rho[j0] * yy, 1e-300), 1.0)
        else:
            gamma = 1.0 / np.maximum(np.linalg.norm(pg, axis=1), 1.0)     # first step: unit-length at most
        r = q * np.asarray(gamma)[:, None]
        for j in reversed(order):
            beta = rho[j] * np.einsum("sd,sd->s", yh[j], r)
            r += (alpha[j] - beta)[:, None] * sh[j]
        dirn = np.where(free, -r, 0.0)
        slope = np.einsum("sd,sd->s", dirn, g)
        fallback = ~(slope < 0.0)
        if fallback.any():
            scale = 1.0 / np.maximum(np.linalg.norm(pg, axis=1), 1.0)
            dirn[fallback] = -(pg * scale[:, None])[fallback]
        # Armijo backtracking along the projection arc, all unfinished starts together
        t = np.ones(S)
        need = ~done
        x_new = x.copy(); f_new = f.copy(); g_new = g.copy()
        for _ in range(max_backtracks):
            xt = np.clip(x + t[:, None] * dirn, lower, upper)
            xt[~need] = x_new[~need]
            ft, gt = fg(xt)
            evals += 1
            ft = np.asarray(ft, float); gt = np.asarray(gt, float)
            dec = np.einsum("sd,sd->s", g, xt - x)
            ok = need & np.isfinite(ft) & np.all(np.isfinite(gt), axis=1) & (ft <= f + c1 * dec)
            x_new[ok] = xt[ok]; f_new[ok] = ft[ok]; g_new[ok] = gt[ok]
            need &= ~ok
            if not need.any():
                break
            t[need] *= 0.5
        stalled = need                                                     # no acceptable step: stop there
        s = x_new - x
        yv = g_new - g
        sy = np.einsum("sd,sd->s", s, yv)
        keep = (~done) & (~stalled) & (sy > 1e-10 * np.linalg.norm(s, axis=1) * np.linalg.norm(yv, axis=1)) & (sy > 0.0)
        slot = it % m
        sh[slot] = np.where(keep[:, None], s, 0.0)
        yh[slot] = np.where(keep[:, None], yv, 0.0)
        rho[slot] = np.where(keep, 1.0 / np.where(keep, sy, 1.0), 0.0)
        small = (f - f_new) <= factr * _EPS * np.maximum(np.maximum(np.abs(f), np.abs(f_new)), 1.0)
        active = ~done
        iters[active] += 1
        x, f, g = x_new, f_new, g_new
        done |= stalled | (active & small)
    return {"x": x, "f": f, "iterations": iters, "evaluations": evals, "converged": done}


CRIT_IDS = {"EI": 0, "quantile": 1, "AKG": 2, "AEI": 3, "EQI": 4, "qEI": 5}


def device_polish(model, criterion, x0, lower, upper, q=1, type="UK", minimization=True, proxy=False, plugin=0.0,
                  qnorm_beta=0.0, new_noise_var=0.0, maxit=100, history=8, pgtol=0.0, factr=1e7, max_backtracks=25, device=0):
    """dk_polish (include/dk_b200.h): the same projected L-BFGS with the optimiser state resident on the device and one
    criterion launch per trip for all starts.  x0: S x D (D = q d; for qEI the rows of a start are its q points, row-major
    like ``x.reshape(S, q, d)``).  Returns dict(x[S, D], f[S] = criterion value at x, iterations[S])."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    d = model.d
    x0 = np.asarray(x0, float).reshape(-1, q * d)
    S = x0.shape[0]
    rows = np.asfortranarray(x0.reshape(S * q, d))                      # (S q) x d column-major
    lower = np.ascontiguousarray(np.broadcast_to(np.asarray(lower, float).ravel()[:d], (d,)))
    upper = np.ascontiguousarray(np.broadcast_to(np.asarray(upper, float).ravel()[:d], (d,)))
    crit = _lib.dk_crit_spec(criterion=CRIT_IDS[criterion], type={"SK": 0, "UK": 1}[type], minimization=int(bool(minimization)),
                             proxy=int(bool(proxy)), plugin=float(plugin), qnorm_beta=float(qnorm_beta),
                             new_noise_var=float(new_noise_var))
    opt = _lib.dk_lbfgs_options(maxit=int(maxit), history=int(history), pgtol=float(pgtol), factr=float(factr),
                                max_backtracks=int(max_backtracks))
    out_x = np.empty((S * q, d), order="F"); out_v = np.empty(S); out_it = np.empty(S, dtype=np.int32)
    vp = lambda a: C.c_void_p(a.ctypes.data)
    _lib.check(lib.dk_polish(model.handle(device).ptr, C.byref(crit), int(q), vp(rows), S, vp(lower), vp(upper), C.byref(opt),
                             vp(out_x), vp(out_v), vp(out_it)), "dk_polish")
    return {"x": np.ascontiguousarray(out_x).reshape(S, q * d), "f": out_v, "iterations": out_it}
