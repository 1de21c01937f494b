"""CPU tests of the oracle itself: the strongest pinning available without an R install
(the reference ships no golden vectors; its self-validating examples are
finite-difference and Monte-Carlo cross-checks, mirrored here)."""
import mpmath
import numpy as np
import pytest

from oracle import dicekriging as dk, diceoptim as do, testfuns
from conftest import small_model, to_oracle


def test_test_functions_match_reference_constants():
    # R/hartman4.R:31, R/branin2.R:32-34, R/rosenbrock4.R:32-34
    assert abs(testfuns.hartman4([0.1873, 0.1906, 0.5566, 0.2647])[0] - (-3.135474)) < 1e-6
    assert abs(testfuns.rosenbrock4([0.4] * 4)[0] - (-1.019701)) < 1e-6
    v = testfuns.branin2([[0.1239, 0.8183], [0.5428, 0.1517], [0.9617, 0.1650]])
    assert np.all(np.abs(v - (-1.047410)) < 2e-5)


def test_pnorm_dnorm_qnorm_against_mpmath():
    mpmath.mp.dps = 40
    for x in [-38.0, -20.0, -8.3, -3.0, -0.5, 0.0, 0.7, 2.5, 6.0, 8.2]:
        ref = float(mpmath.ncdf(x))
        assert do.pnorm(x) == pytest.approx(ref, rel=5e-14, abs=1e-300)
        assert do.dnorm(x) == pytest.approx(float(mpmath.npdf(x)), rel=5e-13, abs=1e-300)
    for p in [1e-10, 0.01, 0.1, 0.5, 0.9, 0.999]:
        assert float(mpmath.ncdf(do.qnorm(p))) == pytest.approx(p, rel=1e-12)


@pytest.mark.parametrize("name", dk.KERNELS)
@pytest.mark.parametrize("trend", ["const", "linear"])
def test_interpolation_and_gradients_by_finite_differences(name, trend):
    m = to_oracle(small_model(name, trend, n=25, d=2, seed=3))
    p = dk.predict_km(m, m.X[:5])
    assert np.allclose(p["mean"], m.y[:5], atol=1e-8)
    assert np.all(p["sd"] < 1e-5)
    rng = np.random.default_rng(5)
    for typ in ("UK", "SK"):
        x = rng.random(2)
        h = 1e-6
        g = do.EI_grad(x, m, type=typ)
        fd = np.array([(do.EI(x + h * e, m, type=typ)[0] - do.EI(x - h * e, m, type=typ)[0]) / (2 * h) for e in np.eye(2)])
        assert np.allclose(g, fd, rtol=2e-5, atol=1e-9)
        gq = do.kriging_quantile_grad(x, m, type=typ)
        fdq = np.array([(do.kriging_quantile(x + h * e, m, type=typ)[0] - do.kriging_quantile(x - h * e, m, type=typ)[0]) / (2 * h)
                        for e in np.eye(2)])
        assert np.allclose(gq, fdq, rtol=2e-5, atol=1e-9)


def test_ei_edge_semantics():
    m = to_oracle(small_model(n=20, d=2))
    ei = do.EI(m.X[:3], m)                       # at design points sd -> 0
    assert np.all((ei >= 0) | np.isnan(ei))
    x = np.random.default_rng(0).random((7, 2))
    assert np.allclose(do.EI(x, m, proxy=True), np.min(m.y) - dk.predict_km(m, x)["mean"])
    assert np.allclose(do.EI(x, m, minimization=False, plugin=-np.max(m.y)), do.EI(x, m, minimization=False))
    assert do.which_max([np.nan, 1.0, 3.0, 3.0, np.nan]) == 2
    assert do.which_max([np.nan, np.nan]) == -1


def test_akg_runs_and_is_nonnegative():
    from diceoptim_b200 import synth
    X = synth.lhs(12, 2, 421)
    rng = np.random.default_rng(421)
    y = synth.branin2(X) + np.sqrt(0.2) * rng.standard_normal(12)
    m = to_oracle(synth.make_km(X, y, "matern5_2", [0.4, 0.5], float(np.var(y)), noise_var=0.2))
    vals = [do.AKG(x, m, new_noise_var=0.2) for x in rng.random((10, 2))]
    assert all(np.isfinite(v) and v >= -1e-12 for v in vals)


def test_akg_grad_oracle_vs_finite_differences_sk():
    """SK only: for UK the reference's mu.x.grad carries a spurious factor 2 (R/AKG.grad.R:148), restated as is."""
    from diceoptim_b200 import synth
    rng = np.random.default_rng(3)
    m = to_oracle(small_model("matern5_2", "const", n=25, d=2, seed=421, noise_var=0.02 + 0.03 * rng.random(25)))
    checked = 0
    for _ in range(12):
        x = rng.random(2)
        g = do.AKG_grad(x, m, new_noise_var=0.02, type="SK")
        if np.max(np.abs(g)) < 1e-9:
            continue
        h = 1e-6
        fd = np.array([(do.AKG(x + h * e, m, new_noise_var=0.02, type="SK") - do.AKG(x - h * e, m, new_noise_var=0.02, type="SK")) / (2 * h)
                       for e in np.eye(2)])
        assert np.max(np.abs(g - fd)) <= 1e-3 * np.max(np.abs(g)) + 1e-12
        checked += 1
    assert checked >= 2


@pytest.mark.parametrize("type", ["SK", "UK"])
def test_akg_against_its_definition_by_monte_carlo(type):
    """AKG(x) = E_Z[max_i (A_i + B_i Z)] - max_i A_i with A = -(kriging means at the design and at x) and
    B = cov_n(., x) / sqrt(tau^2 + s_n(x)^2) (R/AKG.R:136-146).  Here A and B come from predict.km's own posterior
    covariance (cov.compute branch) instead of AKG's V.X / mu.x algebra, and the expectation from 2e6 normal draws
    instead of the envelope scan: an independent check of both."""
    rng = np.random.default_rng(3)
    X = rng.random((15, 2)); y = np.sin(5 * X[:, 0]) + X[:, 1] + 0.1 * rng.standard_normal(15)
    m = dk.make_km(X, y, "matern5_2", np.array([0.4, 0.5]), float(np.var(y)), noise_var=np.full(15, 0.02))
    tau2 = 0.03
    for x in rng.random((3, 2)):
        pts = np.vstack([X, x])
        pr = dk.predict_km(m, pts, type, cov_compute=True)
        A = -pr["mean"]
        sk2 = pr["cov"][-1, -1]
        B = pr["cov"][:, -1] / np.sqrt(tau2 + sk2)
        Z = rng.standard_normal(2_000_000)
        best = np.full(Z.size, -np.inf)
        for a, b in zip(A, B):
            np.maximum(best, a + b * Z, out=best)
        mc = best.mean() - A.max(); se = best.std() / np.sqrt(Z.size)
        got = do.AKG(x, m, new_noise_var=tau2, type=type)
        assert abs(got - mc) < 5 * se + 1e-9, (got, mc, se)


def test_ei_against_its_definition_by_quadrature():
    """EI(x) = E[max(plugin - Y(x), 0)], Y(x) ~ N(mean, sd^2) (minimisation), by adaptive quadrature."""
    from scipy.integrate import quad
    from scipy.stats import norm
    km = to_oracle(small_model("matern5_2", "linear", n=25, d=2, seed=5))
    x = np.random.default_rng(1).random((6, 2))
    pr = dk.predict_km(km, x, "UK")
    plugin = float(np.median(km.y))
    got = do.EI(x, km, plugin=plugin)
    for i in range(6):
        m, s = pr["mean"][i], pr["sd"][i]
        ref = quad(lambda y: (plugin - y) * norm.pdf(y, m, s), m - 12 * s, plugin, epsabs=1e-14, epsrel=1e-12)[0] if plugin > m - 12 * s else 0.0
        assert abs(got[i] - ref) <= 1e-9 * abs(ref) + 1e-13


@pytest.mark.parametrize("type", ["SK", "UK"])
def test_aei_eqi_gradients_by_finite_differences(type):
    """R/AEI.grad.R and R/EQI.grad.R restated vs central differences of the restated AEI / EQI (fixed y.min / q.min)."""
    km = to_oracle(small_model("matern5_2", "linear", n=30, d=2, seed=8, noise_var=0.04))
    rng = np.random.default_rng(6)
    h = 1e-6
    for x in rng.random((3, 2)) * 0.8 + 0.1:
        for f, g, kw in ((do.AEI, do.AEI_grad, dict(new_noise_var=0.03, y_min=float(np.median(km.y)))),
                         (do.EQI, do.EQI_grad, dict(new_noise_var=0.03, beta=0.9, q_min=float(np.median(km.y))))):
            an = g(x, km, type=type, **kw)
            fd = np.array([(float(np.ravel(f(x + h * e, km, type=type, **kw))[0]) - float(np.ravel(f(x - h * e, km, type=type, **kw))[0])) / (2 * h)
                           for e in np.eye(2)])
            assert np.max(np.abs(an - fd)) <= 2e-6 * np.max(np.abs(fd)) + 1e-9, (f.__name__, an, fd)
