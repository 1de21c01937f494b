"""Appending an observation to a km model without re-factorising it.

The reference's constant-liar loop (R/max_qEI.R:305-316) appends the new point to ``model@X`` / ``model@y`` /
``model@F`` and calls DiceKriging's ``computeAuxVariables``, which rebuilds ``T = chol(C)`` (n^3 / 3 flops), ``z`` and
``M`` from scratch with the hyper-parameters and trend coefficients unchanged.  The last column of the new Cholesky
factor only needs one forward substitution against the old one:

    v = T'^-1 c(x),  t = sqrt(k(x, x) - v'v),   T+ = [[T, v], [0, t]],
    z+ = [z, (y - f(x)'beta - v'z) / t],        M+ = [M; (f(x) - v'M) / t]

which is what a column-by-column Cholesky computes for that column, in n^2 flops.  The device handle of the new
model is built on first use like any other.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import solve_triangular

from .km import KM
from .synth import _corr


def update_km(model: KM, x_new, y_new, new_noise_var: float = 0.0, reuse_handles: bool = True) -> KM:
    """The km model with (x_new, y_new) appended; covariance parameters and trend.coef kept (the reference's
    ``computeAuxVariables`` after rbind of X, y, F).  Models with ``noise.var`` give the new point
    ``new_noise_var`` (the reference's loop uses 0 there, R/max_qEI.R:311-314); nugget models add their nugget.
    reuse_handles: True: the device handles of ``model`` are grown in place (dk_model_append: only the last block rows of
    the tiled factor are rebuilt and uploaded) and MOVE to the returned model, ``model`` loses them; "clone": replicas of
    them (device-to-device copies) are grown instead and ``model`` keeps its own; False: the new model builds its handle
    from scratch on first use."""
    x = np.asarray(x_new, float).reshape(1, -1)
    if x.shape[1] != model.d:
        raise ValueError("x_new does not have the right size")
    cov = model.covariance
    theta = np.broadcast_to(np.asarray(cov.range_val, float), (model.d,))
    c = cov.sd2 * _corr(cov.name, model.X, x, theta, cov.shape_val)[:, 0]
    knn = cov.sd2 + (cov.nugget if cov.nugget_flag else (float(new_noise_var) if model.noise_var is not None else 0.0))
    v = solve_triangular(model.T, c, trans="T", lower=False)
    t2 = knn - float(v @ v)
    if not t2 > 0.0:
        raise np.linalg.LinAlgError("the new point makes the covariance matrix numerically singular "
                                    "(checkPredict guards against this, R/checkPredict.R)")
    t = np.sqrt(t2)
    n = model.n
    T = np.zeros((n + 1, n + 1))
    T[:n, :n] = model.T; T[:n, n] = v; T[n, n] = t
    f = np.ones(1) if model.trend == "const" else np.concatenate([[1.0], x[0]])
    M = np.atleast_2d(model.M).reshape(n, model.p)
    z_new = (float(y_new) - float(f @ model.trend_coef) - float(v @ model.z)) / t
    m_new = (f - v @ M) / t
    nv = None if model.noise_var is None else np.append(model.noise_var, float(new_noise_var))
    handles = {}
    if reuse_handles and model._handles:
        from . import _lib
        import ctypes as C
        lib = _lib.load()
        xc = np.ascontiguousarray(x[0])
        for dev, h in list(model._handles.items()):
            if reuse_handles == "clone":
                h = h.clone()
            _lib.check(lib.dk_model_append(h.ptr, C.c_void_p(xc.ctypes.data), float(y_new), float(new_noise_var)),
                       "dk_model_append")
            h.n = n + 1
            handles[dev] = h
        if reuse_handles != "clone":
            model._handles.clear()
    new = KM(X=np.vstack([model.X, x]), y=np.append(model.y, float(y_new)), F=np.vstack([model.F, f[None, :]]),
              trend=model.trend, trend_coef=model.trend_coef, covariance=cov, T=T, z=np.append(model.z, z_new),
              M=np.vstack([M, m_new[None, :]]), noise_var=nv)
    new._handles.update(handles)
    return new
