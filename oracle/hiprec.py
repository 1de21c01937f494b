"""Extended-precision restatement of the predict.km / EI / EI.grad / kriging.quantile(.grad) / AKG arithmetic.

TEST INFRASTRUCTURE ONLY (like the rest of oracle/): imported by tests/ and tools/parity_table.py, never by the
product.  It follows the same reference lines as oracle/dicekriging.py and oracle/diceoptim.py
(/root/reference/R/EI.R:112-132, R/EI.grad.R:149-180, R/kriging.quantile.R:102-105,
R/kriging.quantile.grad.R:116-137, R/AKG.R:105-205, and DiceKriging's predict.km) but carries every
intermediate in numpy's long double (x87 80-bit, eps 1.1e-19) and evaluates pnorm / dnorm with mpmath at 40 digits.
Its inputs are the SAME float64 model slots (T, z, M, beta, theta) the oracle and the GPU get, so its output is
the exact value of the reference's formulas for those inputs to ~1e-17: where the float64 oracle and the GPU
disagree beyond 1e-10, this says which side is off (profiles/parity_r02.md).
"""
from __future__ import annotations

import numpy as np
import mpmath as mp

LD = np.longdouble
mp.mp.dps = 40


def _ld(a):
    return np.asarray(a, dtype=LD)


def _to_mp(x):
    hi = float(x)
    return mp.mpf(hi) + mp.mpf(float(LD(x) - LD(hi)))


def _from_mp(v):
    hi = float(v)
    return LD(hi) + LD(float(v - mp.mpf(hi)))


def pnorm(x):
    return _from_mp(mp.ncdf(_to_mp(x)))


def dnorm(x):
    return _from_mp(mp.npdf(_to_mp(x)))


def _neg_log_corr(h, cov):
    th = _ld(cov.range_val)
    name = cov.name
    if name == "gauss":
        s = LD(0.5) * (h / th) ** 2
    elif name == "exp":
        s = h / th
    elif name == "matern3_2":
        a = np.sqrt(LD(3.0)) * h / th
        s = a - np.log1p(a)
    elif name == "matern5_2":
        a = np.sqrt(LD(5.0)) * h / th
        s = a - np.log(LD(1.0) + a + a * a / LD(3.0))
    elif name == "powexp":
        s = (h / th) ** _ld(cov.shape_val)
    else:
        raise ValueError(name)
    return s.sum(axis=-1)


def cov_mat1_mat2(cov, X1, X2, nugget_flag=False):
    X1 = _ld(np.atleast_2d(X1)); X2 = _ld(np.atleast_2d(X2))
    h = np.abs(X1[:, None, :] - X2[None, :, :])
    C = LD(cov.sd2) * np.exp(-_neg_log_corr(h, cov))
    if nugget_flag and cov.nugget_flag:
        C = C + LD(cov.nugget) * (h.sum(axis=-1) < 1e-15)
    return C


def cov_vector_dx(cov, x, X, c):
    x = _ld(x).ravel(); X = _ld(np.atleast_2d(X))
    diff = x[None, :] - X
    h = np.abs(diff); sg = np.sign(diff); th = _ld(cov.range_val)
    name = cov.name
    if name == "gauss":
        dl = -diff / th ** 2
    elif name == "exp":
        dl = -sg / th
    elif name == "matern3_2":
        a = np.sqrt(LD(3.0)) * h / th
        dl = -sg * (np.sqrt(LD(3.0)) / th) * a / (LD(1.0) + a)
    elif name == "matern5_2":
        a = np.sqrt(LD(5.0)) * h / th
        dl = -sg * (np.sqrt(LD(5.0)) / th) * (a / LD(3.0)) * (LD(1.0) + a) / (LD(1.0) + a + a * a / LD(3.0))
    elif name == "powexp":
        p = _ld(cov.shape_val)
        with np.errstate(divide="ignore", invalid="ignore"):
            dl = -sg * p * np.where(h > 0, h ** (p - LD(1.0)), LD(0.0)) / th ** p
    else:
        raise ValueError(name)
    return _ld(c).reshape(-1, 1) * dl


def forward_solve(T, B):
    """backsolve(t(T), B, upper.tri=FALSE): T upper (float64 values), B n x m; long-double substitution."""
    T = _ld(T); V = _ld(B).copy()
    if V.ndim == 1:
        V = V[:, None]
    n = T.shape[0]
    for i in range(n):
        if i:
            V[i] -= T[:i, i] @ V[:i]
        V[i] /= T[i, i]
    return V


def _chol_upper(G):
    G = _ld(G); p = G.shape[0]
    R = np.zeros((p, p), LD)
    for j in range(p):
        for i in range(j + 1):
            s = G[i, j] - R[:i, i] @ R[:i, j]
            R[i, j] = s / R[i, i] if i < j else np.sqrt(s)
    return R


def _inv_spd(G):
    R = _chol_upper(G); p = R.shape[0]
    Y = forward_solve(R, np.eye(p))          # R' Y = I
    X = np.zeros((p, p), LD)
    for i in range(p - 1, -1, -1):           # R X = Y
        X[i] = (Y[i] - R[i, i + 1:] @ X[i + 1:]) / R[i, i]
    return X


def _model_matrix(trend, X):
    X = _ld(np.atleast_2d(X))
    one = np.ones((X.shape[0], 1), LD)
    return one if trend == "const" else np.hstack([one, X])


def predict_km(model, newdata, type="UK"):
    """predict.km in long double: dict with mean, sd, c, Tinv.c (n x N), S (p x N or None)."""
    newdata = np.atleast_2d(np.asarray(newdata, float))
    cov = model.covariance
    z = _ld(model.z); M = _ld(np.atleast_2d(model.M).reshape(model.n, -1)); beta = _ld(model.trend_coef)
    Fn = _model_matrix(model.trend, newdata)
    c = cov_mat1_mat2(cov, model.X, newdata, nugget_flag=cov.nugget_flag)
    v = forward_solve(model.T, c)
    mean = Fn @ beta + v.T @ z
    total = LD(cov.sd2) + (LD(cov.nugget) if cov.nugget_flag else LD(0.0))
    s2 = total - np.einsum("ij,ij->j", v, v)
    S = None
    if type == "UK":
        TM = _chol_upper(M.T @ M)
        S = forward_solve(TM, (Fn - v.T @ M).T)
        s2 = s2 + np.einsum("ij,ij->j", S, S)
    s2 = np.maximum(s2, LD(0.0))
    return {"mean": mean, "sd": np.sqrt(s2), "c": c, "Tinv.c": v, "S": S}


def back_solve(T, B):
    """backsolve(T, B): T upper (float64 values), long-double substitution from the last row up."""
    T = _ld(T); V = _ld(B).copy()
    if V.ndim == 1:
        V = V[:, None]
    n = T.shape[0]
    for i in range(n - 1, -1, -1):
        if i < n - 1:
            V[i] -= T[i, i + 1:] @ V[i + 1:]
        V[i] /= T[i, i]
    return V


def predict_sensitivity(model, newdata, type="UK"):
    """How far predict.km's mean and sd^2 move when every entry of the covariance vector c(x) is off by a relative delta:
    |d mean| <= delta * sum_i |w_i c_i| with w = T^-1 z, and |d sd^2| <= delta * 2 sum_i |g_i c_i| with
    g = T^-1 (v + M TM^-1 S) (UK) or T^-1 v (SK).  Any float64 evaluation carries delta of a few ulp (exp, the scaled
    distance), so on an ill-conditioned design (w of size 1e6 with alternating signs) this is the error no float64
    implementation can avoid; tests use it as the allowance where the float64 oracle itself misses the tolerance."""
    pr = predict_km(model, newdata, type=type)
    c, v = pr["c"], pr["Tinv.c"]
    w = back_solve(model.T, _ld(model.z))[:, 0]
    g = v
    if type == "UK":
        M = _ld(np.atleast_2d(model.M).reshape(model.n, -1))
        TM = _chol_upper(M.T @ M)
        g = v + M @ back_solve(TM, pr["S"])
    g = back_solve(model.T, g)
    return {"mean": np.abs(w)[:, None].T @ np.abs(c), "sd2": 2.0 * np.einsum("ij,ij->j", np.abs(g), np.abs(c))}


def EI(x, model, plugin=None, type="UK", minimization=True, proxy=False, pred=None):
    """R/EI.R:112-132 over the rows of x."""
    if plugin is None:
        plugin = float(np.min(model.y)) if minimization else -float(np.max(model.y))
    pred = pred or predict_km(model, x, type)
    mean = pred["mean"] if minimization else -pred["mean"]
    sd = pred["sd"]
    out = np.empty(mean.shape, LD)
    for i in range(mean.size):
        diff = LD(plugin) - mean[i]
        if proxy:
            out[i] = diff
            continue
        xcr = diff / sd[i]
        out[i] = diff * pnorm(xcr) + sd[i] * dnorm(xcr)
    return out


def _mean_sd2_grad(x, model, type, v, c):
    cov = model.covariance
    z = _ld(model.z); U = _ld(np.atleast_2d(model.M).reshape(model.n, -1))
    Fn = _model_matrix(model.trend, x[None, :])
    dc = cov_vector_dx(cov, x, model.X, c)
    d = model.d
    fdx = np.zeros((1, d), LD) if model.trend == "const" else np.vstack([np.zeros((1, d), LD), np.eye(d, dtype=LD)])
    W = forward_solve(model.T, dc)
    mean_grad = W.T @ z + _ld(model.trend_coef) @ fdx
    if type == "UK":
        tuuinv = _inv_spd(U.T @ U)
        sd2_grad = (-LD(2.0) * v @ W + LD(2.0) * (Fn - v @ U) @ tuuinv @ (fdx - (W.T @ U).T)).ravel()
    else:
        sd2_grad = (-LD(2.0) * v @ W).ravel()
    return mean_grad, sd2_grad


def EI_grad(x, model, plugin=None, type="UK", minimization=True, proxy=False):
    """R/EI.grad.R:97-186 at one point."""
    if plugin is None:
        plugin = float(np.min(model.y)) if minimization else -float(np.max(model.y))
    x = np.asarray(x, float).ravel()
    pred = predict_km(model, x[None, :], type)
    mean = pred["mean"][0] if minimization else -pred["mean"][0]
    sd = pred["sd"][0]
    if sd / np.sqrt(LD(model.covariance.sd2)) < 1e-6:
        return np.zeros(x.size, LD)
    mean_grad, sd2_grad = _mean_sd2_grad(x, model, type, pred["Tinv.c"][:, 0], pred["c"][:, 0])
    if not minimization:
        mean_grad = -mean_grad
    if proxy:
        return -mean_grad
    xcr = (LD(plugin) - mean) / sd
    return -mean_grad * pnorm(xcr) + sd2_grad / (LD(2.0) * sd) * dnorm(xcr)


def kriging_quantile(x, model, qnorm_beta, type="UK", pred=None):
    pred = pred or predict_km(model, x, type)
    return pred["mean"] + LD(qnorm_beta) * pred["sd"]


def kriging_quantile_grad(x, model, qnorm_beta, type="UK"):
    x = np.asarray(x, float).ravel()
    pred = predict_km(model, x[None, :], type)
    sk = pred["sd"][0]
    mk_grad, sk2_grad = _mean_sd2_grad(x, model, type, pred["Tinv.c"][:, 0], pred["c"][:, 0])
    if sk < np.sqrt(LD(model.covariance.sd2)) / 1e6:
        return mk_grad
    return mk_grad + LD(qnorm_beta) * sk2_grad / (LD(2.0) * sk)


def AKG(x, model, new_noise_var=0.0, type="UK", predX=None):
    """R/AKG.R:93-229 at one point, everything up to the envelope in long double (the envelope scan itself is
    oracle.diceoptim._akg_envelope's control flow on long-double a, b).  predX = predict_km(model, model.X, type)
    may be passed in (candidate independent, R/AKG.R:121)."""
    x = np.asarray(x, float).ravel()
    pred = predict_km(model, x[None, :], type)
    mk_x = pred["mean"][0]; sk_x = pred["sd"][0]
    c_x = pred["c"][:, 0]; V_x = pred["Tinv.c"][:, 0]
    U = _ld(np.atleast_2d(model.M).reshape(model.n, -1))
    sd2 = model.covariance.sd2
    if sk_x < np.sqrt(LD(sd2)) / 1e6 or sd2 < 1e-20:
        return LD(0.0)
    predX = predX or predict_km(model, model.X, type)
    mk_X = predX["mean"]; V_X = predX["Tinv.c"]
    m_min = min(np.min(mk_X), mk_x)
    if type == "UK":
        tuuinv = _inv_spd(U.T @ U)
        F_x = _model_matrix(model.trend, x[None, :])
        mu_x = ((_ld(model.F) - V_X.T @ U) @ tuuinv @ (F_x - V_x[None, :] @ U).T).ravel()
    else:
        mu_x = LD(0.0)
    cn = np.append(c_x - V_X.T @ V_x + mu_x, sk_x ** 2)
    A = -np.append(mk_X, mk_x)
    B = cn / np.sqrt(LD(new_noise_var) + sk_x ** 2)
    return _envelope(A, B, model.n) + m_min


def _envelope(A, B, nobs):
    Isort = np.lexsort((A, B))
    b = B[Isort]; a = A[Isort]
    keep = np.ones(b.size, bool)
    for i in range(nobs):
        if b[i + 1] == b[i]:
            keep[i] = False
    b = b[keep]; a = a[keep]
    nobs = a.size - 1
    C = np.zeros(nobs + 2, LD); C[0] = -1e36; C[-1] = 1e36
    A1 = [0]
    for k in range(2, nobs + 2):
        nondom = 1
        if k != nobs + 1 and a[k] >= a[k - 1] and b[k - 1] == b[k]:
            nondom = 0
        if nondom:
            count = 0
            while count < 1000:
                count += 1
                u = A1[-1] + 1
                C[u] = (a[u - 1] - a[k - 1]) / (b[k - 1] - b[u - 1])
                if len(A1) > 1 and C[u] <= C[A1[-2] + 1]:
                    A1.pop()
                else:
                    A1.append(k - 1)
                    break
    A1 = np.asarray(A1)
    at = a[A1]; bt = b[A1]
    ct = C[np.concatenate(([0], A1 + 1))]
    tot = LD(0.0)
    for k in range(at.size):
        tot += at[k] * (pnorm(ct[k + 1]) - pnorm(ct[k])) + bt[k] * (dnorm(ct[k]) - dnorm(ct[k + 1]))
    return tot
