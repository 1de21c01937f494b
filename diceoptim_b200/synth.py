"""Synthetic workloads of BASELINE.json's configs: designs, objective values and km objects with
FIXED hyper-parameters.  This is input generation, not the hot path: the Cholesky
factor etc. are what DiceKriging's ``km()`` / ``computeAuxVariables`` would hand over; the engine
consumes them.  numpy's default_rng is used, not R's RNG (there is no R on the box).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import cholesky, solve_triangular

from .km import KM, Covariance


def lhs(n: int, d: int, seed: int) -> np.ndarray:
    """Random Latin hypercube in [0,1]^d (same construction as R/sampleFromEI.R:143-149 randLHS)."""
    rng = np.random.default_rng(seed)
    perm = np.argsort(rng.random((n, d)), axis=0)
    return (perm + rng.random((n, d))) / n


def branin2(x):
    x = np.atleast_2d(x)
    xx = 15.0 * x[:, 0] - 5.0; y = 15.0 * x[:, 1]
    f = (y - 5.1 * xx ** 2 / (4 * np.pi ** 2) + 5 * xx / np.pi - 6) ** 2 + 10 * (1 - 1 / (8 * np.pi)) * np.cos(xx) + 10
    return (f - 54.8104) / 51.9496


def hartman4(x):
    x = np.atleast_2d(x)
    a = np.array([[10.0, 3.0, 17.0, 3.5], [0.05, 10.0, 17.0, 0.1], [3.0, 3.5, 1.7, 10.0], [17.0, 8.0, 0.05, 10.0]])
    p = np.array([[0.1312, 0.1696, 0.5569, 0.0124], [0.2329, 0.4135, 0.8307, 0.3736],
                  [0.2348, 0.1451, 0.3522, 0.2883], [0.4047, 0.8828, 0.8732, 0.5743]])
    c = np.array([1.0, 1.2, 3.0, 3.2])
    dd = ((x[:, None, :] - p[None]) ** 2 * a[None]).sum(-1)
    return (-(c[None] * np.exp(-dd)).sum(-1) + 1.1) / 0.8387


def rosenbrock4(x):
    x = np.atleast_2d(x)
    xx = 15.0 * x - 5.0
    f = (100.0 * (xx[:, :-1] ** 2 - xx[:, 1:]) ** 2 + (1 - xx[:, :-1]) ** 2).sum(1)
    return (f - 382658.057227524) / 375264.858362295


def sphere_w(x):
    """'sphere6-style' weighted shifted sphere in any dimension (R/sphere6.R:34-56 without the rotation)."""
    x = np.atleast_2d(x)
    i = np.arange(x.shape[1])
    w = 2.0 ** ((i % 6) + 1); s = 0.5 * np.cos(1.0 + i)
    f = (w * (10.0 * x - 5.0 + s) ** 2).sum(1)
    mu = (w * (100.0 / 12.0 + s ** 2)).sum()
    return (f - mu) / mu


def _corr(name, X1, X2, theta, shape):
    h = np.abs(X1[:, None, :] - X2[None, :, :])
    if name == "gauss":
        s = 0.5 * (h / theta) ** 2
    elif name == "exp":
        s = h / theta
    elif name == "matern3_2":
        a = np.sqrt(3.0) * h / theta; s = a - np.log1p(a)
    elif name == "matern5_2":
        a = np.sqrt(5.0) * h / theta; s = a - np.log(1.0 + a + a * a / 3.0)
    elif name == "powexp":
        s = (h / theta) ** shape
    else:
        raise ValueError(name)
    return np.exp(-s.sum(-1))


def make_km(X, y, name, theta, sd2, trend="const", beta=None, shape=None, nugget=None, noise_var=None) -> KM:
    """A km object with fixed hyper-parameters: C = sd2 R + diag(nugget | noise.var), T = chol(C),
    GLS trend coefficients, z = T'^-1 (y - F beta), M = T'^-1 F."""
    X = np.atleast_2d(np.asarray(X, float)); y = np.asarray(y, float).ravel()
    n, d = X.shape
    theta = np.broadcast_to(np.asarray(theta, float), (d,)).copy()
    shp = None if shape is None else np.broadcast_to(np.asarray(shape, float), (d,)).copy()
    Cm = sd2 * _corr(name, X, X, theta, shp)
    Cm = 0.5 * (Cm + Cm.T); np.fill_diagonal(Cm, sd2)
    nv = None
    if nugget is not None:
        Cm = Cm + nugget * np.eye(n)
    elif noise_var is not None:
        nv = np.broadcast_to(np.asarray(noise_var, float), (n,)).copy()
        Cm = Cm + np.diag(nv)
    T = cholesky(Cm, lower=False)
    F = np.ones((n, 1)) if trend == "const" else np.hstack([np.ones((n, 1)), X])
    M = solve_triangular(T, F, trans="T", lower=False)
    if beta is None:
        beta = np.linalg.lstsq(M, solve_triangular(T, y, trans="T", lower=False), rcond=None)[0]
    beta = np.asarray(beta, float).ravel()
    z = solve_triangular(T, y - F @ beta, trans="T", lower=False)
    cov = Covariance(name=name, sd2=float(sd2), range_val=theta, shape_val=shp,
                     nugget=0.0 if nugget is None else float(nugget), nugget_flag=nugget is not None)
    return KM(X=X, y=y, F=F, trend=trend, trend_coef=beta, covariance=cov, T=T, z=z, M=M, noise_var=nv)


# BASELINE.json configs; hyper-parameters fixed, seeds fixed.
CONFIGS = {
    "cfg1_branin2_n20": dict(d=2, n=20, seed=1, fun=branin2, name="matern5_2", theta=0.3, N=10_000, cand="grid"),
    "cfg2_hartman4_n200": dict(d=4, n=200, seed=2, fun=hartman4, name="gauss", theta=0.35, N=1_000_000, cand_seed=12),
    "cfg3_rosenbrock4_n500": dict(d=4, n=500, seed=3, fun=rosenbrock4, name="matern5_2", theta=0.5, N=400_000, cand_seed=13),
    "cfg4_sphere10_n2000": dict(d=10, n=2000, seed=4, fun=sphere_w, name="matern5_2", theta=1.0, N=10_000_000, cand_seed=14),
    "cfg5_noisy20_n5000": dict(d=20, n=5000, seed=5, fun=sphere_w, name="matern5_2", theta=2.0, N=10_000_000, cand_seed=15,
                               noise=0.1),
}


def config_model(key: str, n: int | None = None, noise_kind: str = "noise_var") -> KM:
    c = CONFIGS[key]
    n = n or c["n"]
    X = lhs(n, c["d"], c["seed"])
    y = c["fun"](X)
    kw = {}
    if "noise" in c:
        y = y + np.sqrt(c["noise"]) * np.random.default_rng(c["seed"] + 100).standard_normal(n)
        kw = {"noise_var": c["noise"]} if noise_kind == "noise_var" else {"nugget": c["noise"]}
    return make_km(X, y, c["name"], c["theta"], float(np.var(y)), trend="const", **kw)


def config_candidates(key: str, N: int | None = None, rank: int = 0) -> np.ndarray:
    """N x d candidates, column-major.  cfg1: 100 x 100 grid in expand.grid order (first coord fastest)."""
    c = CONFIGS[key]
    N = N or c["N"]
    if c.get("cand") == "grid":
        g = int(round(np.sqrt(N)))
        u = np.linspace(0.0, 1.0, g)
        return np.asfortranarray(np.column_stack([np.tile(u, g), np.repeat(u, g)]))
    rng = np.random.default_rng([c["cand_seed"], rank])
    return np.asfortranarray(rng.random((c["d"], N)).T)
