"""Multivariate normal CDFs for q <= 4 — the oracle's stand-in for mnormt::pmnorm / dmnorm.

mnormt is an unpinned third-party dependency of the reference (DESCRIPTION:14; call sites
R/qEI.R:160,166,179, R/qEI.grad.R:4,22,24,349,394) and is not under /root/reference.  For d <= 3 it
wraps Genz's bivariate / trivariate routines, for d > 3 Genz's sadmvn with maxpts = q*200 and
abseps = 1e-6, i.e. the reference's own values carry ~1e-6 absolute error.  This module restates the
published algorithms in deterministic form:

  Phi_1 : scipy ndtr (Cody-class accuracy, like R's pnorm)
  Phi_2 : Genz (2004) BVND, Drezner-Wesolowsky with 20-point Gauss-Legendre          (~1e-15)
  Phi_3 : Plackett's formula as in Genz's TVPACK: Phi(h1) Phi_2(h2,h3) + a path integral over the
          correlations of variable 1, r_1j(t) = sin(t asin r_1j), inner factor Phi_1
  Phi_4 : the same reduction one dimension up: Phi(h1) Phi_3(h2,h3,h4) + path integral whose inner
          factor is the bivariate CDF of the remaining pair conditioned on (x_1, x_j) = (h_1, h_j)

The path integrals use a fixed composite Gauss-Legendre rule (PANELS x NODES; same tables as the CUDA
kernels, tools/gen_gl_tables.py), so the GPU can be compared at 1e-10; `adaptive=True` swaps in
scipy.integrate.quad (epsabs 1e-13) to measure the fixed rule's own error (tests/test_oracle_mvnorm.py).
"""
from __future__ import annotations

import numpy as np
from scipy.integrate import quad
from scipy.special import ndtr

TWOPI = 2.0 * np.pi
PANELS = 2
NODES = 16
_GL20_X, _GL20_W = np.polynomial.legendre.leggauss(20)
_GLN_X, _GLN_W = np.polynomial.legendre.leggauss(NODES)


def phi(x):
    return np.exp(-0.5 * x * x) / np.sqrt(TWOPI)


def bvnu(dh: float, dk: float, r: float) -> float:
    """P(X > dh, Y > dk), corr r: Genz's BVND (20-point rule on every branch)."""
    if dh == np.inf or dk == np.inf:
        return 0.0
    if dh == -np.inf:
        return float(ndtr(-dk))
    if dk == -np.inf:
        return float(ndtr(-dh))
    x = _GL20_X[10:]  # positive half nodes; loop uses +-x
    w = _GL20_W[10:]
    h, k = dh, dk
    hk = h * k
    bvn = 0.0
    if abs(r) < 0.925:
        if abs(r) > 0:
            hs = (h * h + k * k) / 2.0
            asr = np.arcsin(r)
            for xi, wi in zip(x, w):
                for sgn in (-1.0, 1.0):
                    sn = np.sin(asr * (sgn * xi + 1.0) / 2.0)
                    bvn += wi * np.exp((sn * hk - hs) / (1.0 - sn * sn))
            bvn = bvn * asr / (2.0 * TWOPI)
        bvn += ndtr(-h) * ndtr(-k)
    else:
        if r < 0:
            k = -k
            hk = -hk
        if abs(r) < 1:
            as_ = (1.0 - r) * (1.0 + r)
            a = np.sqrt(as_)
            bs = (h - k) ** 2
            c = (4.0 - hk) / 8.0
            d = (12.0 - hk) / 16.0
            asr = -(bs / as_ + hk) / 2.0
            if asr > -100:
                bvn = a * np.exp(asr) * (1.0 - c * (bs - as_) * (1.0 - d * bs / 5.0) / 3.0 + c * d * as_ * as_ / 5.0)
            if -hk < 100:
                b = np.sqrt(bs)
                bvn -= np.exp(-hk / 2.0) * np.sqrt(TWOPI) * ndtr(-b / a) * b * (1.0 - c * bs * (1.0 - d * bs / 5.0) / 3.0)
            a = a / 2.0
            for xi, wi in zip(x, w):
                for sgn in (-1.0, 1.0):
                    xs = (a * (sgn * xi + 1.0)) ** 2
                    rs = np.sqrt(1.0 - xs)
                    asr = -(bs / xs + hk) / 2.0
                    if asr > -100:
                        sp = 1.0 + c * xs * (1.0 + d * xs)
                        ep = np.exp(-hk * xs / (2.0 * (1.0 + rs) ** 2)) / rs
                        bvn += a * wi * np.exp(asr) * (ep - sp)
            bvn = -bvn / TWOPI
        if r > 0:
            bvn += ndtr(-max(h, k))
        else:
            bvn = -bvn
            if k > h:
                if h < 0:
                    bvn += ndtr(k) - ndtr(h)
                else:
                    bvn += ndtr(-h) - ndtr(-k)
    return float(min(1.0, max(0.0, bvn)))


def bvn_cdf(h: float, k: float, r: float) -> float:
    """Phi_2(h, k; r) = P(X <= h, Y <= k)."""
    return bvnu(-h, -k, r)


def _path_integral(f, adaptive: bool) -> float:
    if adaptive:
        return quad(f, 0.0, 1.0, epsabs=1e-13, epsrel=1e-13, limit=200)[0]
    tot = 0.0
    for pnl in range(PANELS):
        lo = pnl / PANELS
        for xi, wi in zip(_GLN_X, _GLN_W):
            t = lo + (xi + 1.0) / (2.0 * PANELS)
            tot += wi / (2.0 * PANELS) * f(t)
    return tot


def _sincs(x: float):
    """sin(x), cos(x)^2 with Genz's series near |x| = pi/2 (TVPACK SINCS)."""
    ee = (np.pi / 2.0 - abs(x)) ** 2
    if ee < 5e-5:
        sx = np.copysign(1.0 - ee * (1.0 - ee / 12.0) / 2.0, x)
        cs = ee * (1.0 - ee * (1.0 - 2.0 * ee / 15.0) / 3.0)
    else:
        sx = np.sin(x)
        cs = 1.0 - sx * sx
    return sx, cs


def tvn_cdf(h, r12: float, r13: float, r23: float, adaptive: bool = False) -> float:
    """Phi_3(h; R) by Plackett's formula along r_1j(t) = sin(t asin r_1j), j = 2, 3.

    d Phi_3 / d r_1j = phi_2(h1, hj; r_1j) Phi_1(conditional of the third variable); the
    1/sqrt(1 - r^2) of phi_2 cancels against d r_1j / d t = a_j cos(t a_j).
    """
    h1, h2, h3 = (float(v) for v in h)
    base = float(ndtr(h1)) * bvn_cdf(h2, h3, r23)
    a2, a3 = np.arcsin(r12), np.arcsin(r13)
    if abs(a2) + abs(a3) == 0.0:
        return base

    def term(hj, hk, rj, cs_j, rk, rjk):
        # variables (1, j) at (h1, hj) with corr rj (cs_j = 1 - rj^2); third variable k: corr rk with 1, rjk with j
        if cs_j <= 0.0:
            return 0.0
        e = np.exp(-(h1 * h1 - 2.0 * rj * h1 * hj + hj * hj) / (2.0 * cs_j))
        mu = ((rk - rj * rjk) * h1 + (rjk - rj * rk) * hj) / cs_j
        var = 1.0 - (rk * rk - 2.0 * rj * rk * rjk + rjk * rjk) / cs_j
        if var <= 0.0:
            return e * (1.0 if hk >= mu else 0.0)
        return e * float(ndtr((hk - mu) / np.sqrt(var)))

    def f(t):
        s2, c2 = _sincs(a2 * t)
        s3, c3 = _sincs(a3 * t)
        v = 0.0
        if a2 != 0.0:
            v += a2 * term(h2, h3, s2, c2, s3, r23)
        if a3 != 0.0:
            v += a3 * term(h3, h2, s3, c3, s2, r23)
        return v

    return float(min(1.0, max(0.0, base + _path_integral(f, adaptive) / TWOPI)))


def qvn_cdf(h, R, adaptive: bool = False) -> float:
    """Phi_4(h; R), R a 4 x 4 correlation matrix: Phi(h1) Phi_3(h2..4) + path integral over r_1j(t),
    j = 2..4, whose inner factor is the bivariate CDF of the other two variables given (x1, xj)."""
    h = np.asarray(h, float)
    R = np.asarray(R, float)
    base = float(ndtr(h[0])) * tvn_cdf(h[1:], R[1, 2], R[1, 3], R[2, 3], adaptive=adaptive)
    a = np.arcsin(R[0, 1:])
    if np.all(a == 0.0):
        return base

    def f(t):
        sc = [_sincs(aj * t) for aj in a]
        r1 = np.array([s for s, _ in sc])           # r_1j(t), j = 2..4
        v = 0.0
        for jj in range(3):
            if a[jj] == 0.0:
                continue
            rj, csj = sc[jj]
            if csj <= 0.0:
                continue
            j = jj + 1
            oth = [m for m in (1, 2, 3) if m != j]
            k, l = oth
            hj = h[j]
            e = np.exp(-(h[0] ** 2 - 2.0 * rj * h[0] * hj + hj * hj) / (2.0 * csj))
            # conditional law of (x_k, x_l) given x_1 = h1, x_j = hj under R(t)
            rk1, rl1 = r1[k - 1], r1[l - 1]
            rkj, rlj, rkl = R[k, j], R[l, j], R[k, l]
            mk = ((rk1 - rj * rkj) * h[0] + (rkj - rj * rk1) * hj) / csj
            ml = ((rl1 - rj * rlj) * h[0] + (rlj - rj * rl1) * hj) / csj
            vk = 1.0 - (rk1 * rk1 - 2.0 * rj * rk1 * rkj + rkj * rkj) / csj
            vl = 1.0 - (rl1 * rl1 - 2.0 * rj * rl1 * rlj + rlj * rlj) / csj
            ckl = rkl - (rk1 * rl1 - rj * (rk1 * rlj + rl1 * rkj) + rkj * rlj) / csj
            if vk <= 0.0 or vl <= 0.0:
                # degenerate conditional: fall back to the product of indicator / univariate pieces
                pk = (1.0 if h[k] >= mk else 0.0) if vk <= 0.0 else float(ndtr((h[k] - mk) / np.sqrt(vk)))
                pl = (1.0 if h[l] >= ml else 0.0) if vl <= 0.0 else float(ndtr((h[l] - ml) / np.sqrt(vl)))
                inner = min(pk, pl) if (vk <= 0.0 and vl <= 0.0) else pk * pl
            else:
                rho = max(-1.0, min(1.0, ckl / np.sqrt(vk * vl)))
                inner = bvn_cdf((h[k] - mk) / np.sqrt(vk), (h[l] - ml) / np.sqrt(vl), rho)
            v += a[jj] * e * inner
        return v

    return float(min(1.0, max(0.0, base + _path_integral(f, adaptive) / TWOPI)))


# ---- dimension >= 5: Genz's separation-of-variables transform on a fixed lattice ----
# mnormt::pmnorm calls Genz's sadmvn for d > 3 (randomised Korobov lattices, maxpts = 200 q in the reference, abseps 1e-6:
# its values carry 1e-4 .. 1e-6 of error).  Deterministic restatement of the same transform (Genz 1992, "Numerical
# computation of multivariate normal probabilities", with the greedy variable reordering of Gibson, Glasbey & Elston):
#   1. standardise; Cholesky factor C of the correlation matrix, choosing at step i the variable with the smallest
#      conditional probability e_j given the expected values y_l of the variables already placed;
#   2. f(w) = prod_i e_i(w),  e_0 = Phi(b_0 / C_00),  y_{i-1} = Phi^-1(w_{i-1} e_{i-1}),
#      e_i = Phi((b_i - sum_{l<i} C_il y_l) / C_ii),  w in [0,1]^(q-1);
#   3. mean of f over LATTICE_N points of the Richtmyer sequence w_kj = |2 frac(k sqrt(p_j)) - 1| (p_j the j-th prime,
#      k = 1..N; the tent map periodises the integrand).
# Error of the rule itself: about 1e-5 (tests/test_oracle_mvnorm.py, against the q <= 4 rules and Monte-Carlo).  The CUDA
# kernels (csrc/dk_mvn_lattice.cuh) run the same steps, so GPU and oracle agree to rounding.
LATTICE_N = 8192
LATTICE_MAX_Q = 10
_SQRT_PRIMES = np.sqrt(np.array([2.0, 3.0, 5.0, 7.0, 11.0, 13.0, 17.0, 19.0, 23.0]))
_ONE_M = 1.0 - 2.0 ** -53
_INV_SQRT_2PI = 0.3989422804014326779


def sov_factor(z, R):
    """Step 1: reordered limits b and Cholesky factor C (lower) of the reordered correlation matrix."""
    from scipy.special import ndtr as _ndtr
    q = z.size
    b = np.array(z, float); R = np.array(R, float); C = np.zeros((q, q)); y = np.zeros(q)
    for i in range(q):
        best, bj = np.inf, i
        for j in range(i, q):
            v = max(R[j, j] - float(C[j, :i] @ C[j, :i]), 1e-280)
            e = float(_ndtr((b[j] - float(C[j, :i] @ y[:i])) / np.sqrt(v)))
            if e < best:
                best, bj = e, j
        if bj != i:
            R[[i, bj]] = R[[bj, i]]; R[:, [i, bj]] = R[:, [bj, i]]
            b[[i, bj]] = b[[bj, i]]; C[[i, bj]] = C[[bj, i]]
        C[i, i] = np.sqrt(max(R[i, i] - float(C[i, :i] @ C[i, :i]), 1e-280))
        for j in range(i + 1, q):
            C[j, i] = (R[j, i] - float(C[j, :i] @ C[i, :i])) / C[i, i]
        t = (b[i] - float(C[i, :i] @ y[:i])) / C[i, i]
        e = float(_ndtr(t))
        y[i] = -_INV_SQRT_2PI * np.exp(-0.5 * t * t) / max(e, 1e-300)
    return b, C


def pmnorm_lattice(z, R, npts: int = LATTICE_N) -> float:
    """Phi_q(z; R) for standardised limits z and correlation R by the fixed-lattice Genz rule above."""
    from scipy.special import ndtri
    z = np.asarray(z, float); q = z.size
    if q > LATTICE_MAX_Q:
        raise ValueError(f"pmnorm: dimension {q} > {LATTICE_MAX_Q}")
    b, C = sov_factor(z, R)
    k = np.arange(1, npts + 1, dtype=float)[:, None]
    t = k * _SQRT_PRIMES[None, :q - 1]
    w = np.abs(2.0 * (t - np.floor(t)) - 1.0)
    e = np.full(npts, float(ndtr(b[0] / C[0, 0])))
    f = e.copy()
    Y = np.zeros((npts, q))
    for i in range(1, q):
        Y[:, i - 1] = ndtri(np.clip(w[:, i - 1] * e, 1e-300, _ONE_M))
        e = ndtr((b[i] - Y[:, :i] @ C[i, :i]) / C[i, i])
        f = f * e
    return float(np.sum(f) / npts)


def pmnorm(x, mean=None, varcov=None, adaptive: bool = False) -> float:
    """mnormt::pmnorm(x, mean, varcov): dimension 1..4 by the deterministic rules above, 5..10 by the fixed-lattice
    Genz rule (covariance, not correlation, input)."""
    x = np.atleast_1d(np.asarray(x, float)).ravel()
    q = x.size
    S = np.atleast_2d(np.asarray(varcov, float))
    if mean is not None:
        x = x - np.atleast_1d(np.asarray(mean, float)).ravel()
    sd = np.sqrt(np.diag(S))
    z = x / sd
    R = S / np.outer(sd, sd)
    if q == 1:
        return float(ndtr(z[0]))
    if q == 2:
        return bvn_cdf(z[0], z[1], max(-1.0, min(1.0, R[0, 1])))
    R = np.clip(R, -1.0, 1.0)
    if q == 3:
        return tvn_cdf(z, R[0, 1], R[0, 2], R[1, 2], adaptive=adaptive)
    if q == 4:
        return qvn_cdf(z, R, adaptive=adaptive)
    return pmnorm_lattice(z, R)


def dmnorm(x, mean, varcov) -> float:
    """mnormt::dmnorm: multivariate normal density."""
    x = np.atleast_1d(np.asarray(x, float)).ravel() - np.atleast_1d(np.asarray(mean, float)).ravel()
    S = np.atleast_2d(np.asarray(varcov, float))
    q = x.size
    sol = np.linalg.solve(S, x)
    return float(np.exp(-0.5 * x @ sol) / np.sqrt((TWOPI ** q) * np.linalg.det(S)))
