"""Restatement of the DiceKriging (>= 1.2; CRAN 1.6.0 assumed) routines that DiceOptim's hot
path calls.  DiceKriging is a third-party dependency that is NOT under /root/reference
(DESCRIPTION:13 only names it), so these functions restate its published algorithm; each cites
the DiceOptim call sites that pin the semantics.  float64 throughout; scipy's
``solve_triangular`` is LAPACK dtrtrs / BLAS dtrsm, the same chain R's ``backsolve`` uses.

Conventions (R): matrices are (rows, cols) numpy arrays; ``T`` is the UPPER Cholesky factor
with C = T'T; ``z = T'^{-1}(y - F beta)``; ``M = T'^{-1}F``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np
from scipy.linalg import cholesky, solve_triangular

KERNELS = ("gauss", "exp", "matern3_2", "matern5_2", "powexp")
WHITE_NOISE_EPS = 1e-15  # DiceKriging C_covWhiteNoise: sum_k |x1_k - x2_k| < 1e-15


@dataclass
class Covariance:
    """covTensorProduct slots read at R/EI.grad.R:122,153, R/AKG.R:115, R/qEI.grad.R:107."""
    name: str
    sd2: float
    range_val: np.ndarray           # theta, length d
    shape_val: Optional[np.ndarray] = None  # powexp exponents, length d
    nugget: float = 0.0
    nugget_flag: bool = False


@dataclass
class KM:
    """The km S4 slots the hot path reads."""
    X: np.ndarray                   # n x d design
    y: np.ndarray                   # n
    F: np.ndarray                   # n x p trend matrix
    trend: str                      # "const" (~1) or "linear" (~.)
    trend_coef: np.ndarray          # p
    covariance: Covariance
    noise_var: Optional[np.ndarray] = None   # n, or None
    T: np.ndarray = field(default=None)      # n x n upper Cholesky
    z: np.ndarray = field(default=None)      # n
    M: np.ndarray = field(default=None)      # n x p

    @property
    def n(self): return self.X.shape[0]
    @property
    def d(self): return self.X.shape[1]
    @property
    def p(self): return self.F.shape[1]
    @property
    def noise_flag(self): return self.noise_var is not None


# ----------------------------------------------------------------------------------------
# covariance kernels
# ----------------------------------------------------------------------------------------
def _neg_log_corr(h: np.ndarray, cov: Covariance) -> np.ndarray:
    """Sum over dimensions of -log k1d(h_k); h has shape (..., d), h >= 0.

    DiceKriging's C kernels accumulate ``s`` over dimensions and return sd2*exp(-s)
    (SURVEY 3.6): gauss (h/t)^2/2, exp h/t, matern3_2 a-log(1+a), matern5_2
    a-log(1+a+a^2/3), powexp (h/t)^p.
    """
    th = cov.range_val
    name = cov.name
    if name == "gauss":
        s = 0.5 * (h / th) ** 2
    elif name == "exp":
        s = h / th
    elif name == "matern3_2":
        a = np.sqrt(3.0) * h / th
        s = a - np.log1p(a)
    elif name == "matern5_2":
        a = np.sqrt(5.0) * h / th
        s = a - np.log(1.0 + a + a * a / 3.0)
    elif name == "powexp":
        s = (h / th) ** cov.shape_val
    else:
        raise ValueError(f"unsupported covariance kernel {name!r}")
    return s.sum(axis=-1)


def cov_mat1_mat2(cov: Covariance, X1: np.ndarray, X2: np.ndarray, nugget_flag: bool = False) -> np.ndarray:
    """covMat1Mat2(object, X1, X2, nugget.flag): n1 x n2 cross-covariance.

    Called (through predict.km) from R/EI.R:113,115; directly at R/checkPredict.R:56.
    The nugget is added only where rows coincide and only when BOTH the argument flag and
    covariance@nugget.flag are set (never for noise.var models).
    """
    X1 = np.atleast_2d(X1); X2 = np.atleast_2d(X2)
    h = np.abs(X1[:, None, :] - X2[None, :, :])
    C = cov.sd2 * np.exp(-_neg_log_corr(h, cov))
    if nugget_flag and cov.nugget_flag:
        C = C + cov.nugget * (h.sum(axis=-1) < WHITE_NOISE_EPS)
    return C


def cov_matrix(cov: Covariance, X: np.ndarray, noise_var: Optional[np.ndarray] = None) -> np.ndarray:
    """covMatrix(object, X, noise.var)[[1]] (call sites R/qEI.grad.R:114; predict cov.compute)."""
    X = np.atleast_2d(X)
    C = cov_mat1_mat2(cov, X, X, nugget_flag=False)
    C = 0.5 * (C + C.T)
    np.fill_diagonal(C, cov.sd2)
    if cov.nugget_flag:
        C = C + cov.nugget * np.eye(X.shape[0])
    elif noise_var is not None:
        C = C + np.diag(np.asarray(noise_var, float))
    return C


def cov_vector_dx(cov: Covariance, x: np.ndarray, X: np.ndarray, c: np.ndarray) -> np.ndarray:
    """covVector.dx(object, x, X, c): n x d matrix c_i * d log k / d x_k.

    Call sites R/EI.grad.R:157, R/kriging.quantile.grad.R:116, R/AKG.grad.R:134,
    R/qEI.grad.R:118,126.
    """
    x = np.asarray(x, float).ravel(); X = np.atleast_2d(X)
    diff = x[None, :] - X
    h = np.abs(diff); sg = np.sign(diff); th = cov.range_val
    name = cov.name
    if name == "gauss":
        dl = -diff / th ** 2
    elif name == "exp":
        dl = -sg / th
    elif name == "matern3_2":
        a = np.sqrt(3.0) * h / th
        dl = -sg * (np.sqrt(3.0) / th) * a / (1.0 + a)
    elif name == "matern5_2":
        a = np.sqrt(5.0) * h / th
        dl = -sg * (np.sqrt(5.0) / th) * (a / 3.0) * (1.0 + a) / (1.0 + a + a * a / 3.0)
    elif name == "powexp":
        p = cov.shape_val
        with np.errstate(divide="ignore", invalid="ignore"):
            dl = -sg * p * np.where(h > 0, h ** (p - 1.0), 0.0) / th ** p
    else:
        raise ValueError(name)
    return np.asarray(c, float).reshape(-1, 1) * dl


# ----------------------------------------------------------------------------------------
# trend
# ----------------------------------------------------------------------------------------
def model_matrix(trend: str, X: np.ndarray) -> np.ndarray:
    """model.matrix(trend.formula, data) for ~1 and ~. (R/EI.grad.R:149, R/AKG.R:113)."""
    X = np.atleast_2d(X)
    if trend == "const":
        return np.ones((X.shape[0], 1))
    if trend == "linear":
        return np.hstack([np.ones((X.shape[0], 1)), X])
    raise ValueError(f"unsupported trend {trend!r}")


def trend_deltax(trend: str, d: int) -> np.ndarray:
    """trend.deltax(x, model): p x d Jacobian (R/EI.grad.R:158). zeros for ~1, rbind(0, I) for ~."""
    if trend == "const":
        return np.zeros((1, d))
    if trend == "linear":
        return np.vstack([np.zeros((1, d)), np.eye(d)])
    raise ValueError(trend)


# ----------------------------------------------------------------------------------------
# model set-up (computeAuxVariables; fitting itself is out of scope: hyper-parameters fixed)
# ----------------------------------------------------------------------------------------
def make_km(X, y, name, theta, sd2, trend="const", beta=None, shape=None, nugget=None,
            noise_var=None) -> KM:
    """km with FIXED hyper-parameters + computeAuxVariables (call sites R/max_qEI.R:298,303).

    C = sd2 R(X,X) + diag(nugget | noise.var); T = chol(C) upper; beta by GLS when not given;
    z = T'^{-1}(y - F beta); M = T'^{-1}F.
    """
    X = np.atleast_2d(np.asarray(X, float)); y = np.asarray(y, float).ravel()
    d = X.shape[1]
    theta = np.broadcast_to(np.asarray(theta, float), (d,)).copy()
    shp = None if shape is None else np.broadcast_to(np.asarray(shape, float), (d,)).copy()
    cov = Covariance(name=name, sd2=float(sd2), range_val=theta, shape_val=shp,
                     nugget=0.0 if nugget is None else float(nugget), nugget_flag=nugget is not None)
    nv = None if noise_var is None else np.broadcast_to(np.asarray(noise_var, float), y.shape).copy()
    F = model_matrix(trend, X)
    C = cov_matrix(cov, X, noise_var=nv)
    T = cholesky(C, lower=False)
    M = solve_triangular(T, F, trans="T", lower=False)
    if beta is None:
        yt = solve_triangular(T, y, trans="T", lower=False)
        beta = np.linalg.lstsq(M, yt, rcond=None)[0]
    beta = np.asarray(beta, float).ravel()
    z = solve_triangular(T, y - F @ beta, trans="T", lower=False)
    return KM(X=X, y=y, F=F, trend=trend, trend_coef=beta, covariance=cov, noise_var=nv, T=T, z=z, M=M)


# ----------------------------------------------------------------------------------------
# predict.km
# ----------------------------------------------------------------------------------------
def predict_km(model: KM, newdata: np.ndarray, type: str = "UK", se_compute: bool = True,
               cov_compute: bool = False, light_return: bool = False) -> dict:
    """predict.km (call sites R/EI.R:113,115, R/AKG.R:105,121, R/qEI.R:136)."""
    newdata = np.atleast_2d(np.asarray(newdata, float))
    cov = model.covariance
    T, z, M, beta = model.T, model.z, model.M, model.trend_coef
    Fn = model_matrix(model.trend, newdata)
    trend = Fn @ beta
    c = cov_mat1_mat2(cov, model.X, newdata, nugget_flag=cov.nugget_flag)      # n x N
    Tinv_c = solve_triangular(T, c, trans="T", lower=False)                      # backsolve(t(T), c, upper.tri=FALSE)
    mean = trend + Tinv_c.T @ z
    out = {"trend": trend, "mean": mean}
    if not light_return:
        out["c"] = c; out["Tinv.c"] = Tinv_c
    S = None
    if type == "UK" and (se_compute or cov_compute):
        TM = cholesky(M.T @ M, lower=False)
        S = solve_triangular(TM, (Fn - Tinv_c.T @ M).T, trans="T", lower=False)  # p x N
    if se_compute:
        total_sd2 = cov.sd2 + (cov.nugget if cov.nugget_flag else 0.0)
        s2_1 = np.einsum("ij,ij->j", Tinv_c, Tinv_c)
        if type == "SK":
            s2 = np.maximum(total_sd2 - s2_1, 0.0)
        elif type == "UK":
            s2_2 = np.einsum("ij,ij->j", S, S)
            s2 = np.maximum(total_sd2 - s2_1 + s2_2, 0.0)
        else:
            raise ValueError(type)
        out["sd"] = np.sqrt(s2)
    if cov_compute:
        Cn = cov_matrix(cov, newdata)
        cc = Cn - Tinv_c.T @ Tinv_c
        if type == "UK":
            cc = cc + S.T @ S
        out["cov"] = cc
    return out
