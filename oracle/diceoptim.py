"""Restatement of DiceOptim's acquisition criteria (the hot path), following
/root/reference/R/*.R line by line, including guards and quirks.  numpy float64.

pnorm/dnorm/qnorm are R's nmath (Cody 1969 / AS241): scipy.special.ndtr / ndtri agree with them
to a few ulp (cross-checked against mpmath in tests/test_oracle_basic.py).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import solve_triangular
from scipy.special import ndtr, ndtri

from . import dicekriging as dk

_INV_SQRT_2PI = 0.3989422804014326779399460599343818684758586311649


def pnorm(x):
    return ndtr(np.asarray(x, float))


def dnorm(x):
    x = np.asarray(x, float)
    return _INV_SQRT_2PI * np.exp(-0.5 * x * x)


def qnorm(p):
    return ndtri(p)


def _default_plugin(model, minimization):
    # R/EI.R:82-88
    return float(np.min(model.y)) if minimization else float(-np.max(model.y))


def EI(x, model, plugin=None, type="UK", minimization=True, envir=None, proxy=False):
    """R/EI.R:80-145.  x: vector (d) or N x d matrix; returns N values (IEEE semantics at sd=0)."""
    if plugin is None:
        plugin = _default_plugin(model, minimization)
    m = plugin
    x = np.atleast_2d(np.asarray(x, float))
    if x.shape[1] != model.d:
        raise ValueError("x does not have the right size")
    predx = dk.predict_km(model, x, type=type, light_return=envir is None)
    mean = predx["mean"]
    if not minimization:
        mean = -mean
    sd = predx["sd"]
    if proxy:
        xcr = prob = dens = None
        res = m - mean
    else:
        with np.errstate(divide="ignore", invalid="ignore"):
            xcr = (m - mean) / sd
            prob = pnorm(xcr)
            dens = dnorm(xcr)
            res = (m - mean) * prob + sd * dens
    if envir is not None:
        envir.update({"xcr": xcr, "xcr.prob": prob, "xcr.dens": dens, "kriging.sd": sd,
                      "c": predx["c"], "Tinv.c": predx["Tinv.c"]})
    return res


def _mean_sd2_grad(x, model, type, v, c, sd_guard_zero):
    """Shared W algebra of R/EI.grad.R:157-175 and R/kriging.quantile.grad.R:116-131."""
    cov = model.covariance
    T, z, u = model.T, model.z, model.M
    Fn = dk.model_matrix(model.trend, x[None, :])
    dc = dk.cov_vector_dx(cov, x, model.X, c)
    fdx = dk.trend_deltax(model.trend, model.d)
    W = solve_triangular(T, dc, trans="T", lower=False)
    mean_grad = W.T @ z + (model.trend_coef @ fdx)
    if sd_guard_zero:
        return mean_grad, None
    if type == "UK":
        tuuinv = np.linalg.inv(u.T @ u)
        sd2_grad = (-2.0 * v @ W + 2.0 * (Fn - v @ u) @ tuuinv @ (fdx - (W.T @ u).T)).ravel()
    else:
        sd2_grad = (-2.0 * v @ W).ravel()
    return mean_grad, sd2_grad


def EI_grad(x, model, plugin=None, type="UK", minimization=True, envir=None, proxy=False):
    """R/EI.grad.R:97-186 at ONE point x (length d)."""
    if plugin is None:
        plugin = _default_plugin(model, minimization)
    m = plugin
    x = np.asarray(x, float).ravel()
    d = x.size
    if d != model.d:
        raise ValueError("x does not have the right size")
    predx = dk.predict_km(model, x[None, :], type=type)
    mean = predx["mean"][0]
    if not minimization:
        mean = -mean
    sd = predx["sd"][0]
    v = predx["Tinv.c"][:, 0]; c = predx["c"][:, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        xcr = (m - mean) / sd
    prob = pnorm(xcr); dens = dnorm(xcr)
    if sd / np.sqrt(model.covariance.sd2) < 1e-6:          # R/EI.grad.R:153
        return np.zeros(d)
    mean_grad, sd2_grad = _mean_sd2_grad(x, model, type, v, c, sd_guard_zero=proxy)
    if not minimization:
        mean_grad = -mean_grad
    if proxy:
        return -mean_grad
    sd_grad = sd2_grad / (2.0 * sd)
    return -mean_grad * prob + sd_grad * dens


def kriging_quantile(x, model, beta=0.1, type="UK", envir=None):
    """R/kriging.quantile.R:92-115 (vectorised over rows like predict)."""
    x = np.atleast_2d(np.asarray(x, float))
    predx = dk.predict_km(model, x, type=type)
    return predx["mean"] + qnorm(beta) * predx["sd"]


def kriging_quantile_grad(x, model, beta=0.1, type="UK", envir=None):
    """R/kriging.quantile.grad.R:83-138 at one point."""
    x = np.asarray(x, float).ravel()
    predx = dk.predict_km(model, x[None, :], type=type)
    sk = predx["sd"][0]
    v = predx["Tinv.c"][:, 0]; c = predx["c"][:, 0]
    guard = sk < np.sqrt(model.covariance.sd2) / 1e6      # :123
    mk_grad, sk2_grad = _mean_sd2_grad(x, model, type, v, c, sd_guard_zero=guard)
    sk_grad = 0.0 if guard else sk2_grad / (2.0 * sk)
    return mk_grad + qnorm(beta) * sk_grad


def AKG(x, model, new_noise_var=0.0, type="UK", envir=None, predX=None):
    """R/AKG.R:93-229 at ONE point (approximate knowledge gradient, written for maximisation).
    predX: optional cache of predict_km(model, model.X, type), the candidate-independent prediction of :121 that the
    reference recomputes on every call; passing it changes no arithmetic (used by tests at n = 5000)."""
    x = np.asarray(x, float).ravel()
    if x.size != model.d:
        raise ValueError("x does not have the right size")
    tau2_new = new_noise_var
    predx = dk.predict_km(model, x[None, :], type=type)
    mk_x = predx["mean"][0]; sk_x = predx["sd"][0]
    c_x = predx["c"][:, 0]; V_x = predx["Tinv.c"][:, 0]
    U = model.M
    F_x = dk.model_matrix(model.trend, x[None, :])
    sd2 = model.covariance.sd2
    if sk_x < np.sqrt(sd2) / 1e6 or sd2 < 1e-20:           # :115
        return 0.0
    if predX is None:
        predX = dk.predict_km(model, model.X, type=type)   # :121 (candidate independent)
    mk_X = predX["mean"]; V_X = predX["Tinv.c"]
    F_X = model.F
    m_min = min(np.min(mk_X), mk_x)
    if type == "UK":
        tuuinv = np.linalg.inv(U.T @ U)
        mu_x = ((F_X - V_X.T @ U) @ tuuinv @ (F_x - V_x[None, :] @ U).T).ravel()
    else:
        mu_x = 0.0
    cn = c_x - V_X.T @ V_x + mu_x
    cn = np.append(cn, sk_x ** 2)
    A = np.append(mk_X, mk_x)
    B = cn / np.sqrt(tau2_new + sk_x ** 2)
    A = -A
    return _akg_envelope(A, B, model.n) - (-m_min)


def _akg_envelope(A, B, nobs):
    """R/AKG.R:147-205: order(B, A), drop equal slopes, upper-envelope scan, Phi/phi sum."""
    Isort = np.lexsort((A, B))                               # order(x=B, y=A)
    b = B[Isort]; a = A[Isort]
    keep = np.ones(b.size, bool)
    for i in range(nobs):                                    # :153-157 (1-based i -> b[i+1]==b[i])
        if b[i + 1] == b[i]:
            keep[i] = False
    b = b[keep]; a = a[keep]
    nobs = a.size - 1
    C = np.zeros(nobs + 2); C[0] = -1e36; C[-1] = 1e36
    A1 = [0]
    for k in range(2, nobs + 2):                             # R: k in 2:(nobs+1), 1-based
        nondom = 1
        if k == nobs + 1:
            nondom = 1
        elif a[k] >= a[k - 1] and b[k - 1] == b[k]:          # a[k+1]>=a[k] && b[k]==b[k+1]
            nondom = 0
        if nondom == 1:
            loopdone = 0; count = 0
            while loopdone == 0 and count < 1000:
                count += 1
                u = A1[-1] + 1                                # 1-based index into a/b
                C[u] = (a[u - 1] - a[k - 1]) / (b[k - 1] - b[u - 1])   # C[u+1] 1-based
                if len(A1) > 1 and C[u] <= C[A1[-2] + 1]:      # C[A1[len-1]+2] 1-based
                    A1.pop()
                else:
                    A1.append(k - 1)
                    loopdone = 1
    A1 = np.asarray(A1)
    at = a[A1]; bt = b[A1]
    ct = C[np.concatenate(([0], A1 + 1))]
    maxNew = 0.0
    for k in range(at.size):
        maxNew += at[k] * (pnorm(ct[k + 1]) - pnorm(ct[k])) + bt[k] * (dnorm(ct[k]) - dnorm(ct[k + 1]))
    return float(maxNew)


def which_max(v):
    """R's which.max: first maximal index, NaN skipped; -1 if nothing comparable (R: integer(0))."""
    v = np.asarray(v, float)
    ok = ~np.isnan(v)
    if not ok.any():
        return -1
    return int(np.flatnonzero(ok & (v == np.max(v[ok])))[0])


def max_crit_screen(x, model, type="UK", minimization=True, proxy=False):
    """The batched screen of R/max_crit.R:104-110: EI over the N x d matrix, which.max, value, row."""
    ei = EI(x, model, plugin=None, type=type, minimization=minimization, envir=None, proxy=proxy)
    I = which_max(ei)
    return I, ei[I], np.atleast_2d(x)[I]


def sample_from_EI_weights(x, model, minimization=True, T=None):
    """The EI weights of R/sampleFromEI.R:109-132 (before the weighted sample())."""
    x = np.atleast_2d(np.asarray(x, float))
    pred = dk.predict_km(model, x, type="UK", light_return=True)
    mn, sn = pred["mean"], pred["sd"]
    if T is None:
        T = np.min(model.y) if minimization else np.max(model.y)
    with np.errstate(divide="ignore", invalid="ignore"):
        uu = (mn - T) / sn
        if minimization:
            uu = -uu
        ei = sn * (uu * pnorm(uu) + dnorm(uu))
    nan = np.isnan(ei)
    ei[nan] = (T - mn)[nan] if minimization else (mn - T)[nan]
    ei = ei * (ei > 0)
    return ei


# ----------------------------------------------------------------------------------------
# qEI (R/qEI.R:119-261)
# ----------------------------------------------------------------------------------------
def covZk(sigma, index):
    """R/qEI.R:240-261: covariance of Z^(k) = (Y_k - Y_1, ..., -Y_k, ..., Y_k - Y_q)."""
    q = sigma.shape[0]
    res = sigma.copy()
    for i in range(q):
        if i != index:
            res[index, i] = res[i, index] = sigma[index, index] - sigma[index, i]
    for i in range(q):
        for j in range(i, q):
            if i != index and j != index:
                res[i, j] = res[j, i] = sigma[i, j] + sigma[index, index] - sigma[index, i] - sigma[index, j]
    return res


def get_sigmaik(sigma, i):
    """R/qEI.R:196-214."""
    q = sigma.shape[0]
    res = sigma - np.outer(sigma[:, i], sigma[:, i]) / sigma[i, i]
    keep = [u for u in range(q) if u != i]
    return res[np.ix_(keep, keep)]


def get_cik(b, m, sigma, i):
    """R/qEI.R:217-227."""
    sigmai = sigma[i, :] / sigma[i, i]
    res = (b - m) - (b[i] - m[i]) * sigmai
    return np.delete(res, i)


def r_round8(x):
    """round(x, digits = 8) as used by the duplicate test R/qEI.R:126 (ties are immaterial here)."""
    return np.round(np.asarray(x, float) * 1e8) / 1e8


def qEI_dedup(x, model):
    """R/qEI.R:124-128: rows of x that survive unique(round(rbind(X, x), 8)), in order."""
    xr = r_round8(x)
    seen = {tuple(r) for r in r_round8(model.X)}
    keep = []
    for j, row in enumerate(map(tuple, xr)):
        if row not in seen:
            seen.add(row)
            keep.append(j)
    return keep


def qEI(x, model, plugin=None, type="UK", minimization=True, fastCompute=True, eps=1e-5, envir=None,
        adaptive=False):
    """R/qEI.R:119-194.  fastCompute=TRUE is the reference's finite-difference form of the second term,
    fastCompute=FALSE the exact Chevalier-Ginsbourger double sum; both use oracle.mvnorm.pmnorm."""
    from . import mvnorm as mv
    d = model.d
    x = np.asarray(x, float).reshape(-1, d)
    keep = qEI_dedup(x, model)
    if len(keep) == 0:
        return 0.0
    x = r_round8(x[keep])                                   # the reference continues with the ROUNDED rows
    if plugin is None:
        plugin = float(np.min(model.y)) if minimization else float(np.max(model.y))   # :129-130 (max, not -max)
    if x.shape[0] == 1:
        return float(EI(x, model, plugin=plugin, type=type, envir=envir)[0])          # :132-134 (minimization default)
    krig = dk.predict_km(model, x, type=type, se_compute=False, cov_compute=True)
    sigma, mu = krig["cov"], krig["mean"]
    q = mu.size
    pk = np.zeros(q); first = np.zeros(q); second = np.zeros(q)
    sym = np.zeros((q, q)); nonsym = np.zeros((q, q))
    for k in range(q):
        Sk = covZk(sigma, k)
        mu_k = mu - mu[k]; mu_k[k] = -mu[k]
        if minimization:
            mu_k = -mu_k
        b_k = np.zeros(q); b_k[k] = -plugin
        if minimization:
            b_k = -b_k
        pk[k] = mv.pmnorm(b_k - mu_k, varcov=Sk, adaptive=adaptive)
        first[k] = (mu[k] - plugin) * pk[k]
        if minimization:
            first[k] = -first[k]
        if fastCompute:
            second[k] = (mv.pmnorm(b_k + eps * Sk[:, k] - mu_k, varcov=Sk, adaptive=adaptive) - pk[k]) / eps
        else:
            for i in range(q):
                nonsym[k, i] = Sk[i, k]
                if i >= k:
                    phi_ik = dnorm((b_k[i] - mu_k[i]) / np.sqrt(Sk[i, i])) / np.sqrt(Sk[i, i])
                    cik = get_cik(b_k, mu_k, Sk, i)
                    sik = get_sigmaik(Sk, i)
                    sym[k, i] = phi_ik * mv.pmnorm(cik, varcov=sik, adaptive=adaptive)
    if not fastCompute:
        sym = sym + sym.T
        sym[np.diag_indices(q)] *= 0.5
        second = np.sum(sym * nonsym)
    if envir is not None:
        envir.update({"pk": pk, "kriging.mean": mu, "kriging.cov": sigma})
        if not fastCompute:
            envir["symetric_term"] = sym
    return float(np.sum(first) + np.sum(second))


def qEI_monte_carlo(x, model, plugin=None, type="UK", minimization=True, nsim=300000, seed=0):
    """The Monte-Carlo cross-check of the reference's example (R/qEI.R:89-104): mean of the
    improvement of the batch minimum over draws of the conditional Gaussian vector."""
    x = np.asarray(x, float).reshape(-1, model.d)
    if plugin is None:
        plugin = float(np.min(model.y)) if minimization else float(np.max(model.y))
    krig = dk.predict_km(model, x, type=type, se_compute=False, cov_compute=True)
    rng = np.random.default_rng(seed)
    L = np.linalg.cholesky(krig["cov"] + 1e-14 * np.eye(x.shape[0]))
    Y = krig["mean"][None, :] + rng.standard_normal((nsim, x.shape[0])) @ L.T
    imp = np.maximum(plugin - Y.min(1), 0.0) if minimization else np.maximum(Y.max(1) - plugin, 0.0)
    return float(imp.mean()), float(imp.std() / np.sqrt(nsim))


# ----------------------------------------------------------------------------------------
# qEI.grad (R/qEI.grad.R:1-142, 267-424)
# ----------------------------------------------------------------------------------------
def _decompo(x, mu, Sigma, k, adaptive=False):
    """R/qEI.grad.R:1-26: density of the components k times the CDF of the rest given them."""
    from . import mvnorm as mv
    x = np.asarray(x, float).ravel(); mu = np.asarray(mu, float).ravel()
    k = list(np.atleast_1d(k))
    q = x.size
    if len(k) == q:
        return mv.dmnorm(x, mu, Sigma)
    rest = [u for u in range(q) if u not in k]
    x1 = x[k]; x2 = x[rest]; mu1 = mu[k]
    S22 = Sigma[np.ix_(rest, rest)]; S21 = Sigma[np.ix_(rest, k)]; S11 = Sigma[np.ix_(k, k)]
    S11inv = np.linalg.inv(S11)
    varcov = S22 - S21 @ S11inv @ S21.T
    varcov = 0.5 * (varcov + varcov.T)
    dg = np.diag(varcov).copy()
    if dg.min() <= 0:
        varcov[np.diag_indices_from(varcov)] = dg * (dg >= 0) + 1e-9
    moy = mu[rest] + S21 @ S11inv @ (x1 - mu1)
    return mv.dmnorm(x1, mu1, S11) * mv.pmnorm(x2, mean=moy, varcov=varcov, adaptive=adaptive)


def GPhi(x, mu, Sigma, adaptive=False):
    """R/qEI.grad.R:29-36: gradient of the q-variate normal CDF."""
    q = np.asarray(x).size
    return np.array([_decompo(x, mu, Sigma, [i], adaptive) for i in range(q)])


def HPhi(x, mu, Sigma, gradient=None, adaptive=False):
    """R/qEI.grad.R:39-60: Hessian of the q-variate normal CDF."""
    x = np.asarray(x, float).ravel(); mu = np.asarray(mu, float).ravel()
    Sigma = np.atleast_2d(np.asarray(Sigma, float))
    q = x.size
    if q == 1:
        s = np.sqrt(Sigma[0, 0])
        return np.array([[-(x[0] - mu[0]) / Sigma[0, 0] * dnorm((x[0] - mu[0]) / s) / s]])
    res = np.zeros((q, q))
    for i in range(q - 1):
        for j in range(i + 1, q):
            res[i, j] = _decompo(x, mu, Sigma, [i, j], adaptive)
    res = res + res.T
    if gradient is None:
        gradient = GPhi(x, mu, Sigma, adaptive)
    res = res - np.diag(((x - mu) * np.asarray(gradient).ravel() + np.diag(Sigma @ res)) / np.diag(Sigma))
    return res


def kriging_deriv(x, model, type="UK"):
    """R/qEI.grad.R:62-142 (krigingDeriv): derivatives of the conditional mean and covariance of a batch.
    mean_jacob[l, j, k] = d m_k / d x_lj ; cov_jacob[l, j, k, i] = d Sigma_ki / d x_lj."""
    x = np.asarray(x, float)
    q, d = x.shape
    T, z, u, cov = model.T, model.z, model.M, model.covariance
    Fn = dk.model_matrix(model.trend, x)
    pred = dk.predict_km(model, x, type=type, se_compute=False, cov_compute=True)
    mean, kcov, v, c = pred["mean"], pred["cov"], pred["Tinv.c"], pred["c"]
    if np.prod(np.diag(kcov) / cov.sd2 < 1e-8):              # :107 (all of them tiny)
        return None
    mj = np.zeros((q, d, q)); cj = np.zeros((q, d, q, q))
    tuuinv = np.linalg.inv(u.T @ u)
    covM = dk.cov_matrix(cov, x)
    fdx = dk.trend_deltax(model.trend, d)
    for k in range(q):
        dc = dk.cov_vector_dx(cov, x[k], model.X, c[:, k])
        W = solve_triangular(T, dc, trans="T", lower=False)
        mj[k, :, k] = z @ W + model.trend_coef @ fdx
        for l in range(q):
            kg = dk.cov_vector_dx(cov, x[k], x[l][None, :], np.array([covM[k, l]])).ravel()
            cj[k, :, k, l] = kg - v[:, l] @ W
            if type == "UK":
                cj[k, :, k, l] += ((fdx.T - W.T @ u) @ tuuinv @ (Fn[l] - v[:, l] @ u)).ravel()
    for l in range(q):
        for j in range(d):
            cj[l, j] = cj[l, j] + cj[l, j].T
    return mean, mj, kcov, cj


def qEI_grad(x, model, plugin=None, type="UK", minimization=True, fastCompute=True, eps=1e-6, envir=None,
             adaptive=False):
    """R/qEI.grad.R:267-424.  Returns the q x d gradient matrix."""
    from . import mvnorm as mv
    if not minimization and plugin is None:
        plugin = -float(np.max(model.y))
    d = model.d
    x = np.asarray(x, float).reshape(-1, d)
    q = x.shape[0]
    xb = np.vstack([model.X, x])
    if len({tuple(r) for r in r_round8(xb)}) != xb.shape[0]:   # :293-295 any duplicate -> zero matrix
        return np.zeros((q, d))
    if plugin is None:
        plugin = float(np.min(model.y))
    if q == 1:
        if not minimization:
            raise ValueError("qEI.grad doesn't work in 'minimization = FALSE' when dim = 1 (in progress).")
        return EI_grad(x[0], model, plugin=plugin, type=type)[None, :]
    kd = kriging_deriv(x, model, type=type)
    if kd is None:
        return np.zeros((q, d))
    kmean, mj, kcov, cj = kd
    if not minimization:
        kmean = -kmean; mj = -mj
    G = np.zeros((q, d))
    b = np.zeros(q)
    L = -np.eye(q)
    for k in range(q):
        bk = b.copy(); bk[k] = plugin
        Lk = L.copy(); Lk[:, k] = 1.0
        mk = Lk @ kmean
        Sigk = Lk @ kcov @ Lk.T
        Sigk = 0.5 * (Sigk + Sigk.T)
        pk = mv.pmnorm(bk, mean=mk, varcov=Sigk, adaptive=adaptive)
        G[k] -= mj[k, :, k] * pk                                            # term A1
        gradpk = GPhi(bk - mk, b, Sigk, adaptive)
        hesspk = HPhi(bk, mk, Sigk, gradpk, adaptive)
        Sdx = np.einsum("ab,ljbc,dc->ljad", Lk, cj, Lk)                      # Lk %*% cov.jacob[l,j,,] %*% t(Lk)
        mdx = np.einsum("ab,ljb->lja", Lk, mj)
        Dpk = 0.5 * np.einsum("ab,ljab->lj", hesspk, Sdx) - np.einsum("a,lja->lj", gradpk, mdx)
        G += (plugin - kmean[k]) * Dpk                                      # term A2
        if fastCompute:
            gradpk1 = GPhi(bk - mk + Sigk[:, k] * eps, b, Sigk, adaptive)
            hesspk1 = HPhi(bk + Sigk[:, k] * eps, mk, Sigk, gradpk1, adaptive)
            f1 = (-np.einsum("lja,a->lj", mdx, gradpk1) + eps * np.einsum("lja,a->lj", Sdx[:, :, :, k], gradpk1)
                  + 0.5 * np.einsum("ljab,ab->lj", Sdx, hesspk1))
            f = -np.einsum("lja,a->lj", mdx, gradpk) + 0.5 * np.einsum("ljab,ab->lj", Sdx, hesspk)
            G += (f1 - f) / eps
        else:
            B = np.zeros((q, d))
            for i in range(q):
                oth = [u for u in range(q) if u != i]
                S_ik, S_ii, mk_i, bk_i = Sigk[i, k], Sigk[i, i], mk[i], bk[i]
                S_mi = Sigk[oth, i]
                ck_pi = bk[oth] - mk[oth] - (bk_i - mk_i) / S_ii * S_mi
                A = Sigk[np.ix_(oth, oth)] - np.outer(S_mi, S_mi) / S_ii
                S_pi = 0.5 * (A + A.T)
                phi_ik = dnorm((bk_i - mk_i) / np.sqrt(S_ii)) / np.sqrt(S_ii)
                dphi_dS = ((bk_i - mk_i) ** 2 / (2 * S_ii ** 2) - 0.5 / S_ii) * phi_ik
                dphi_dm = (bk_i - mk_i) / S_ii * phi_ik
                Phi_ik = mv.pmnorm(ck_pi, varcov=S_pi, adaptive=adaptive)
                G_ik = GPhi(ck_pi, np.zeros(q - 1), S_pi, adaptive)
                H_ik = HPhi(ck_pi, np.zeros(q - 1), S_pi, G_ik, adaptive)
                for l in range(q):
                    for j in range(d):
                        sdx = Sdx[l, j]
                        B1 = sdx[i, k] * phi_ik * Phi_ik
                        B2 = S_ik * (mdx[l, j, i] * dphi_dm + dphi_dS * sdx[i, i]) * Phi_ik
                        dck = (-mdx[l, j, oth] + (mdx[l, j, i] * S_ii + (bk_i - mk_i) * sdx[i, i]) / S_ii ** 2 * S_mi
                               - (bk_i - mk_i) / S_ii * sdx[oth, i])
                        cross = np.outer(sdx[oth, i], S_mi)
                        dS = sdx[np.ix_(oth, oth)] + sdx[i, i] / S_ii ** 2 * np.outer(S_mi, S_mi) - (cross + cross.T) / S_ii
                        B3 = S_ik * phi_ik * (G_ik @ dck + 0.5 * np.sum(H_ik * dS))
                        B[l, j] += B1 + B2 + B3
            G += B
    if np.isnan(G.sum()):
        return np.zeros((q, d))
    return G


# ----------------------------------------------------------------------------------------
# AKG.grad (R/AKG.grad.R:65-281), envir = NULL path
# ----------------------------------------------------------------------------------------
def AKG_grad(x, model, new_noise_var=0.0, type="UK", envir=None):
    """Line-by-line restatement, reference quirks included: mu.x.grad carries a factor 2 (:148), ct.grad lacks
    the 1/(bt[i+1]-bt[i]) scaling (:256; it multiplies a term that vanishes at the breakpoints), and m_min.grad
    is never taken because which.min(c(m_min, mk.x)) is 1 whenever mk.x attains the minimum (:268-276)."""
    x = np.asarray(x, float).ravel()
    d, n = model.d, model.n
    tau2 = new_noise_var
    T, z, U, F_X = model.T, model.z, model.M, model.F
    cov = model.covariance
    predx = dk.predict_km(model, x[None, :], type=type)
    mk_x = predx["mean"][0]; sk_x = predx["sd"][0]
    c_x = predx["c"][:, 0]; V_x = predx["Tinv.c"][:, 0]
    F_x = dk.model_matrix(model.trend, x[None, :])
    if sk_x < np.sqrt(cov.sd2) / 1e6 or cov.sd2 < 1e-20:      # :126
        return np.zeros(d)
    predX = dk.predict_km(model, model.X, type=type)
    mk_X = predX["mean"]; V_X = predX["Tinv.c"]
    if type == "UK":
        tuuinv = np.linalg.inv(U.T @ U)
        mu_x = ((F_X - V_X.T @ U) @ tuuinv @ (F_x - V_x[None, :] @ U).T).ravel()
    else:
        mu_x = 0.0
    cn = c_x - V_X.T @ V_x + mu_x
    cn = np.append(cn, sk_x ** 2)
    a = np.append(mk_X, mk_x)
    den2 = tau2 + sk_x ** 2
    b = cn / np.sqrt(den2)
    sQ = b[n]
    dc = dk.cov_vector_dx(cov, x, model.X, c_x)
    fdx = dk.trend_deltax(model.trend, d)
    W = solve_triangular(T, dc, trans="T", lower=False)
    a_grad = np.zeros((d, n + 1))
    a_grad[:, n] = W.T @ z + model.trend_coef @ fdx
    b_grad = np.zeros((d, n + 1))
    if type == "UK":
        sk2_grad = (-2.0 * V_x @ W + 2.0 * (F_x - V_x[None, :] @ U) @ tuuinv @ (fdx - (W.T @ U).T)).ravel()
        mu_x_grad = (2.0 * (F_X - V_X.T @ U) @ tuuinv @ (fdx - (W.T @ U).T)).T      # d x n
    else:
        sk2_grad = (-2.0 * V_x @ W).ravel()
        mu_x_grad = np.zeros((d, n))
    cn_grad = dc.T - (V_X.T @ W).T + mu_x_grad
    b_grad[:, :n] = cn_grad / np.sqrt(den2) - np.outer(sk2_grad, cn[:n]) * (0.5 / den2 ** 1.5)
    b_grad[:, n] = (sk_x ** 2 * (2 * tau2 + sk_x ** 2) / den2 ** 2) * sk2_grad / (2.0 * sQ)
    A = -a; B = b; B_grad = b_grad; A_grad = -a_grad
    Isort = np.lexsort((A, B))
    bs = B[Isort]; as_ = A[Isort]; bg = B_grad[:, Isort]; ag = A_grad[:, Isort]
    keep = np.ones(bs.size, bool)
    for i in range(n):
        if bs[i + 1] == bs[i]:
            keep[i] = False
    bs = bs[keep]; as_ = as_[keep]; bg = bg[:, keep]; ag = ag[:, keep]
    nobs = as_.size - 1
    C = np.zeros(nobs + 2); C[0] = -1e36; C[-1] = 1e36
    A1 = [0]
    for k in range(2, nobs + 2):
        nondom = 1
        if k != nobs + 1 and as_[k] >= as_[k - 1] and bs[k - 1] == bs[k]:
            nondom = 0
        if nondom:
            count = 0
            while count < 1000:
                count += 1
                u = A1[-1] + 1
                C[u] = (as_[u - 1] - as_[k - 1]) / (bs[k - 1] - bs[u - 1])
                if len(A1) > 1 and C[u] <= C[A1[-2] + 1]:
                    A1.pop()
                else:
                    A1.append(k - 1)
                    break
    A1 = np.asarray(A1)
    at = as_[A1]; bt = bs[A1]; ct = C[np.concatenate(([0], A1 + 1))]
    nt = at.size
    atg = ag[:, A1]; btg = bg[:, A1]
    ctg = np.zeros((d, nt + 1))
    for i in range(nt - 1):
        ctg[:, i + 1] = (atg[:, i] - atg[:, i + 1]) - (at[i] - at[i + 1]) * (btg[:, i + 1] - btg[:, i]) / (bt[i + 1] - bt[i])
    g = np.zeros(d)
    for k in range(nt):
        g += (atg[:, k] * (pnorm(ct[k + 1]) - pnorm(ct[k])) + btg[:, k] * (dnorm(ct[k]) - dnorm(ct[k + 1]))
              + ctg[:, k + 1] * dnorm(ct[k + 1]) * (at[k] + bt[k] * ct[k + 1]) - ctg[:, k] * dnorm(ct[k]) * (at[k] + bt[k] * ct[k]))
    return g


def AEI(x, model, new_noise_var=0.0, y_min=None, type="UK", envir=None):
    """R/AEI.R:105-146 at one point."""
    x = np.asarray(x, float).ravel()
    if y_min is None:
        pr = dk.predict_km(model, model.X, type=type)
        y_min = pr["mean"][int(np.argmin(pr["mean"] + qnorm(0.75) * pr["sd"]))]
    pr = dk.predict_km(model, x[None, :], type=type)
    mk, sk = pr["mean"][0], pr["sd"][0]
    if sk < np.sqrt(model.covariance.sd2) / 1e6:
        return 0.0
    xcr = (y_min - mk) / sk
    return float(((y_min - mk) * pnorm(xcr) + sk * dnorm(xcr)) * (1.0 - np.sqrt(new_noise_var) / np.sqrt(new_noise_var + sk ** 2)))


def EQI(x, model, new_noise_var=0.0, beta=0.9, q_min=None, type="UK", envir=None):
    """R/EQI.R:100-151 at one point."""
    x = np.asarray(x, float).ravel()
    if q_min is None:
        pr = dk.predict_km(model, model.X, type=type)
        q_min = np.min(pr["mean"] + qnorm(beta) * pr["sd"])
    pr = dk.predict_km(model, x[None, :], type=type)
    mk, sk = pr["mean"][0], pr["sd"][0]
    if sk < np.sqrt(model.covariance.sd2) / 1e6:
        return 0.0
    mq = mk + qnorm(beta) * np.sqrt((new_noise_var * sk ** 2) / (new_noise_var + sk ** 2))
    sq = sk ** 2 / np.sqrt(new_noise_var + sk ** 2)
    xcr = (q_min - mq) / sq
    return float(sq * (xcr * pnorm(xcr) + dnorm(xcr)))


def AEI_grad(x, model, new_noise_var=0.0, y_min=None, type="UK", envir=None):
    """R/AEI.grad.R:65-154 at one point."""
    x = np.asarray(x, float).ravel()
    if y_min is None:
        pr = dk.predict_km(model, model.X, type=type)
        y_min = pr["mean"][int(np.argmin(pr["mean"] + qnorm(0.75) * pr["sd"]))]
    pr = dk.predict_km(model, x[None, :], type=type)
    mk, sk = pr["mean"][0], pr["sd"][0]
    if sk < np.sqrt(model.covariance.sd2) / 1e6:
        return np.zeros(model.d)
    xcr = (y_min - mk) / sk; prob = pnorm(xcr); dens = dnorm(xcr)
    ei_val = (y_min - mk) * prob + sk * dens
    pen = 1.0 - np.sqrt(new_noise_var) / np.sqrt(new_noise_var + sk ** 2)
    mk_grad, sk2_grad = _mean_sd2_grad(x, model, type, pr["Tinv.c"][:, 0], pr["c"][:, 0], sd_guard_zero=False)
    sk_grad = sk2_grad / (2.0 * sk)
    ei_grad = -mk_grad * prob + sk_grad * dens
    pen_grad = np.sqrt(new_noise_var) / 2.0 * sk2_grad * (new_noise_var + sk ** 2) ** (-1.5)
    return ei_grad * pen + ei_val * pen_grad


def EQI_grad(x, model, new_noise_var=0.0, beta=0.9, q_min=None, type="UK", envir=None):
    """R/EQI.grad.R:66-146 at one point."""
    x = np.asarray(x, float).ravel()
    if q_min is None:
        pr = dk.predict_km(model, model.X, type=type)
        q_min = np.min(pr["mean"] + qnorm(beta) * pr["sd"])
    pr = dk.predict_km(model, x[None, :], type=type)
    mk, sk = pr["mean"][0], pr["sd"][0]
    if sk < np.sqrt(model.covariance.sd2) / 1e6:
        return np.zeros(model.d)
    t2 = new_noise_var
    mq = mk + qnorm(beta) * np.sqrt(t2 * sk ** 2 / (t2 + sk ** 2))
    sq = sk ** 2 / np.sqrt(t2 + sk ** 2)
    xcr = (q_min - mq) / sq; prob = pnorm(xcr); dens = dnorm(xcr)
    mk_grad, sk2_grad = _mean_sd2_grad(x, model, type, pr["Tinv.c"][:, 0], pr["c"][:, 0], sd_guard_zero=False)
    mq_grad = mk_grad + qnorm(beta) * t2 ** 1.5 / (2.0 * sk * (sk ** 2 + t2) ** 1.5) * sk2_grad
    sq2_grad = (sk ** 2) * (2 * t2 + sk ** 2) / (t2 + sk ** 2) ** 2 * sk2_grad
    sq_grad = sq2_grad / (2.0 * sq)
    return -prob * mq_grad + dens * sq_grad


# ---------------------------------------------------------------------------------------
# Constrained criteria: checkPredict, test_feas_vec, Feas, crit_EFI, prob.of.feasible.improvement
# ---------------------------------------------------------------------------------------
def check_predict_distance(x, models, distance="covdist", type="UK"):
    """`mindist` of R/checkPredict.R:35-66, one value per row of x (the reference is written for one row; the
    euclidean branch recycles x against the design, so it is restated row by row)."""
    x = np.atleast_2d(np.asarray(x, float))
    if distance == "euclidean":
        X = models[0].X
        return np.array([np.min(np.sqrt(np.sum((X - xi) ** 2, axis=1))) for xi in x])          # :35-38
    mindist = np.full(x.shape[0], np.inf)
    for m in models:
        sd2 = m.covariance.sd2
        if distance == "covratio":
            mindist = np.minimum(mindist, dk.predict_km(m, x, type, light_return=True)["sd"] / np.sqrt(sd2))  # :45-47
            continue
        kn_xx = dk.predict_km(m, x, type)["sd"] ** 2                                               # :54
        kn_yy = dk.predict_km(m, m.X, type)["sd"] ** 2                                             # :55
        kxy = dk.cov_mat1_mat2(m.covariance, x, m.X)                                               # :56
        kyy = dk.cov_matrix(m.covariance, m.X)                                                     # :57
        Kinv = np.linalg.inv(m.T.T @ m.T)                                                          # chol2inv(T)
        kn_xy = kxy - kxy @ Kinv @ kyy                                                             # :58
        val = np.sqrt(np.maximum(0.0, kn_xx - 2.0 * np.max(kn_xy - kn_yy[None, :], axis=1)) / sd2)  # :61-62
        mindist = np.minimum(mindist, val)
    return mindist


def test_feas_vec(cst, equality=False, tolConstraints=None):
    """R/test_feas_vec.R:11-36."""
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


test_feas_vec.__test__ = False


def Feas(model, newdata, equality=False, threshold=0.0):
    """R/crit_EFI.R:231-238 (predicts with "UK")."""
    p = dk.predict_km(model, newdata, "UK", light_return=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        hi = np.where(p["sd"] > 0, pnorm((threshold - p["mean"]) / p["sd"]), (threshold >= p["mean"]) * 1.0)
        if not equality:
            return hi
        lo = np.where(p["sd"] > 0, pnorm((-threshold - p["mean"]) / p["sd"]), (-threshold >= p["mean"]) * 1.0)
    return hi - lo


def _eq_tol(nc, equality, cc):
    eqv = np.full(nc, False) if (np.ndim(equality) == 0 and not equality) else np.broadcast_to(np.asarray(equality, bool), (nc,))
    if cc.get("tolConstraints") is None:
        tol = np.zeros(nc); tol[eqv] = 0.05
    else:
        tol = np.broadcast_to(np.asarray(cc["tolConstraints"], float), (nc,))
    return eqv, tol


def crit_EFI(x, model_fun, model_constraint, equality=False, critcontrol=None, plugin=None, type="UK"):
    """R/crit_EFI.R:125-228 for ONE point x."""
    cc = dict(critcontrol or {})
    cst = list(model_constraint)
    nc = len(cst)
    observations = np.column_stack([m.y for m in cst])
    if plugin is None:
        feas = test_feas_vec(observations, equality=equality, tolConstraints=cc.get("tolConstraints"))
        plugin = float(np.min(model_fun.y[feas])) if feas.any() else np.inf
    d = check_predict_distance(x, cst + [model_fun], distance=cc.get("distance") or "covdist", type=type)
    if d[0] < (cc.get("threshold") or 1e-4):
        return 0.0
    Ei = 1.0 if plugin == np.inf else float(EI(x, model_fun, plugin=plugin, type=type)[0])
    eqv, tol = _eq_tol(nc, equality, cc)
    Pf = 1.0
    for i in range(nc):
        Pf *= float(Feas(cst[i], np.atleast_2d(x), equality=bool(eqv[i]), threshold=float(tol[i]))[0])
    return Pf * Ei


def prob_of_feasible_improvement(integration_points, model_fun, model_constraint, equality=False, critcontrol=None,
                                 type="UK"):
    """R/prob.of.feasible.improvement.R:1-81."""
    cc = dict(critcontrol or {})
    cst = list(model_constraint)
    nc = len(cst)
    P = np.atleast_2d(np.asarray(integration_points, float))
    pc = [dk.predict_km(m, P, type) for m in cst]
    po = dk.predict_km(model_fun, P, type)
    eqv, tol = _eq_tol(nc, equality, cc)
    observations = np.column_stack([m.y for m in cst])
    feas = test_feas_vec(observations, equality=equality, tolConstraints=cc.get("tolConstraints"))
    obj_min = float(np.min(model_fun.y[feas])) if feas.any() else np.inf
    with np.errstate(divide="ignore", invalid="ignore"):
        p_obj = pnorm((obj_min - po["mean"]) / po["sd"]) if obj_min != np.inf else np.ones(P.shape[0])
        pf = np.ones(P.shape[0])
        for i in range(nc):
            if eqv[i]:
                pf = pf * (pnorm((tol[i] - pc[i]["mean"]) / pc[i]["sd"]) - pnorm((-tol[i] - pc[i]["mean"]) / pc[i]["sd"]))
            else:
                pf = pf * pnorm((0.0 - pc[i]["mean"]) / pc[i]["sd"])
    return p_obj * pf


# ---------------------------------------------------------------------------------------
# Stepwise uncertainty reduction with constraints: precomputeUpdateData, computeQuickKrigcov2, crit_SUR_cst
# ---------------------------------------------------------------------------------------
def precompute_update_data(model, integration_points):
    """R/precomputeUpdateData.R:1-22."""
    P = np.atleast_2d(np.asarray(integration_points, float))
    c_old = dk.cov_mat1_mat2(model.covariance, model.X, P, nugget_flag=model.covariance.nugget_flag)       # n x M
    Tinv_c = solve_triangular(model.T, c_old, trans="T", lower=False)                                       # the big backsolve
    Kinv_c = solve_triangular(model.T, Tinv_c, lower=False)
    Kinv_F = solve_triangular(model.T, model.M, lower=False)
    inv_tF_Kinv_F = np.linalg.inv(model.M.T @ model.M)                                                      # chol2inv(chol(crossprod(M)))
    fP = dk.model_matrix(model.trend, P)
    first_member = (fP - c_old.T @ Kinv_F) @ inv_tF_Kinv_F
    return {"Kinv.c.olddata": Kinv_c, "Kinv.F": Kinv_F, "first.member": first_member}


def compute_quick_krigcov2(model, integration_points, X_new, precalc, F_newdata, c_newdata):
    """R/computeQuickKrigcov2.R:22-33: kriging covariances between X.new and the integration points (M x q)."""
    P = np.atleast_2d(np.asarray(integration_points, float))
    c_xnew = dk.cov_mat1_mat2(model.covariance, P, np.atleast_2d(X_new), nugget_flag=model.covariance.nugget_flag)
    cov_std = c_xnew - precalc["Kinv.c.olddata"].T @ c_newdata
    if F_newdata is None:
        return cov_std
    second = (F_newdata - c_newdata.T @ precalc["Kinv.F"]).T
    return precalc["first.member"] @ second + cov_std


def _pbivnorm(x, y, rho):
    """pbivnorm::pbivnorm with the device's deterministic BVND (oracle/mvnorm.py); rho clamped to [-1, 1] where
    rounding pushed it outside (the package would stop())."""
    from . import mvnorm as mv
    x, y, rho = np.broadcast_arrays(np.asarray(x, float), np.asarray(y, float), np.asarray(rho, float))
    out = np.empty(x.shape)
    for idx in np.ndindex(x.shape):
        r = rho[idx]
        out[idx] = np.nan if (np.isnan(r) or np.isnan(x[idx]) or np.isnan(y[idx])) else \
            mv.bvn_cdf(float(x[idx]), float(y[idx]), float(min(1.0, max(-1.0, r))))
    return out


def crit_SUR_cst(x, model_fun, model_constraint, equality=False, critcontrol=None, type="UK"):
    """R/crit_SUR_cst.R:120-378 for ONE point x; critcontrol$integration.points must be given
    (integration_design_cst is outside the hot path)."""
    cc = dict(critcontrol or {})
    cst = list(model_constraint)
    nc = len(cst)
    if nc > 3:
        return np.nan                                                                                      # :124-127
    x = np.atleast_2d(np.asarray(x, float))
    d = check_predict_distance(x, cst + [model_fun], distance=cc.get("distance") or "covdist", type=type)
    if d[0] < (cc.get("threshold") or 1e-4):
        return 0.0                                                                                         # :133-136
    P = np.atleast_2d(np.asarray(cc["integration.points"], float))
    Mi = P.shape[0]
    w = cc.get("integration.weights")
    w = np.full(Mi, 1.0 / Mi) if w is None else np.asarray(w, float)
    pre_c, mn_c, sn_c = [], np.zeros((nc, Mi)), np.zeros((nc, Mi))
    for i, m in enumerate(cst):                                                                            # :155-160
        pt = dk.predict_km(m, P, type)
        mn_c[i], sn_c[i] = pt["mean"], pt["sd"]
        pre_c.append(precompute_update_data(m, P))
    pt = dk.predict_km(model_fun, P, type)
    mn_o, sn_o = pt["mean"], pt["sd"]
    pre_o = precompute_update_data(model_fun, P)
    obs = np.column_stack([m.y for m in cst])
    feas = test_feas_vec(obs, equality=equality, tolConstraints=cc.get("tolConstraints"))
    obj_min = float(np.min(model_fun.y[feas])) if feas.any() else np.inf                                   # :178-184
    A = np.array([np.sqrt(m.covariance.sd2) / 1e4 for m in cst])[:, None]
    keep = ~np.all(sn_c < A, axis=0)                                                                       # :192-208
    P, mn_o, sn_o, mn_c, sn_c = P[keep], mn_o[keep], sn_o[keep], mn_c[:, keep], sn_c[:, keep]
    if w.size > 1:
        w = w[keep]
    pre_o = {**pre_o, "Kinv.c.olddata": pre_o["Kinv.c.olddata"][:, keep], "first.member": pre_o["first.member"][keep]}
    pre_c = [{**q, "Kinv.c.olddata": q["Kinv.c.olddata"][:, keep], "first.member": q["first.member"][keep]} for q in pre_c]
    Mi = P.shape[0]
    if Mi == 0:
        return 0.0
    eqv, tol = _eq_tol(nc, equality, cc)
    any_eq = bool(np.any(eqv))
    with np.errstate(divide="ignore", invalid="ignore"):
        ko = dk.predict_km(model_fun, x, type)                                                             # :226-228
        mk, sk = ko["mean"][0], ko["sd"][0]
        kn = compute_quick_krigcov2(model_fun, P, x, pre_o, dk.model_matrix(model_fun.trend, x), ko["c"])[:, 0]
        rho_o = kn / (sk * sn_o)
        den = np.sqrt(sk ** 2 + sn_o ** 2 - 2 * kn)
        eta = (mk - mn_o) / den
        nu = (kn - sk ** 2) / sk / den
        omt = (obj_min - mn_o) / sn_o
        p_omt = pnorm(omt)
        omb = (obj_min - mk) / sk
        w1 = np.zeros(Mi); w2 = np.zeros(Mi)
        if omb > -10:                                                                                      # :243-250
            if omb == np.inf:
                w1 = pnorm(eta)
            else:
                J = np.where(eta > -10)[0]
                w1[J] = _pbivnorm(np.full(J.size, omb), eta[J], nu[J])
        if -omb > -10:                                                                                     # :251-258
            if -omb == np.inf:
                w2 = p_omt
            else:
                J = np.where(omt > -10)[0]
                w2[J] = _pbivnorm(np.full(J.size, -omb), omt[J], -rho_o[J])
        pf_minus, pf_plus = w1 + w2, p_omt
        Tbar = np.zeros(nc); Tbarm = np.zeros(nc); Tbarp = np.zeros(nc)
        Tt = np.zeros((nc, Mi)); Ttm = np.zeros((nc, Mi)); Ttp = np.zeros((nc, Mi)); rho_c = np.zeros((nc, Mi))
        for i, m in enumerate(cst):                                                                        # :274-299
            kc = dk.predict_km(m, x, type)
            mkc, skc = kc["mean"][0], kc["sd"][0]
            knc = compute_quick_krigcov2(m, P, x, pre_c[i], dk.model_matrix(m.trend, x), kc["c"])[:, 0]
            if any_eq and eqv[i]:
                Tbarp[i] = (tol[i] - mkc) / skc; Ttp[i] = (tol[i] - mn_c[i]) / sn_c[i]
                Tbarm[i] = (-tol[i] - mkc) / skc; Ttm[i] = (-tol[i] - mn_c[i]) / sn_c[i]
            else:
                Tbar[i] = (-tol[i] - mkc) / skc; Tt[i] = (-tol[i] - mn_c[i]) / sn_c[i]
            rho_c[i] = knc / (skc * sn_c[i])
        pf = np.ones(Mi)                                                                                   # :303-315
        for i in range(nc):
            pf = pf * ((pnorm(Ttp[i]) - pnorm(Ttm[i])) if (any_eq and eqv[i]) else pnorm(Tt[i]))
        oldcrit = p_omt * pf
        pgm = np.zeros((nc, Mi)); pgp = np.zeros((nc, Mi))
        for i in range(nc):                                                                                # :319-348
            if any_eq and eqv[i]:
                pgm[i] = (_pbivnorm(Tbarp[i], Ttp[i], rho_c[i]) - _pbivnorm(Tbarp[i], Ttm[i], rho_c[i])
                          - _pbivnorm(Tbarm[i], Ttp[i], rho_c[i]) + _pbivnorm(Tbarm[i], Ttm[i], rho_c[i]))
                pgp[i] = pnorm(Ttp[i]) - pnorm(Ttm[i]) - pgm[i]
            else:
                if Tbar[i] > -10:
                    J = np.where(Tt[i] > -10)[0]
                    pgm[i, J] = _pbivnorm(np.full(J.size, Tbar[i]), Tt[i, J], rho_c[i, J])
                pgp[i] = pnorm(Tt[i]) - pgm[i]
        PT = pnorm(Tt)
        if nc == 1:                                                                                        # :351-369
            newcrit = pf_minus * pgm[0] + pf_plus * pgp[0]
        elif nc == 2:
            newcrit = pf_minus * pgm[0] * pgm[1] + pf_plus * (PT[0] * pgp[1] + PT[1] * pgp[0] - pgp[0] * pgp[1])
        else:
            newcrit = pf_minus * pgm[0] * pgm[1] * pgm[2] + pf_plus * (
                PT[1] * PT[2] * pgp[0] + PT[0] * PT[2] * pgp[1] + PT[0] * PT[1] * pgp[2]
                - PT[2] * pgp[0] * pgp[1] - PT[1] * pgp[0] * pgp[2] - PT[0] * pgp[1] * pgp[2] + pgp[0] * pgp[1] * pgp[2])
        a = (oldcrit - newcrit) * w
    ok = ~np.isnan(a)
    return float(np.sum(a[ok]) / np.sum(ok)) if ok.any() else np.nan                                       # mean(., na.rm = TRUE)
