"""CPU: the oracle's deterministic normal CDFs against independent references."""
import numpy as np
from scipy.stats import multivariate_normal as mvn

from oracle import mvnorm as mv
from oracle import dicekriging as dk, diceoptim as do
from conftest import small_model, to_oracle


def _corr(rng, q, s):
    A = rng.standard_normal((q, q + 1)); A[:, 0] *= s
    S = A @ A.T; sd = np.sqrt(np.diag(S))
    return S / np.outer(sd, sd)


def test_bvn_matches_scipy_bivariate():
    rng = np.random.default_rng(0)
    for _ in range(100):
        r = float(np.tanh(rng.standard_normal() * 1.5)); h, k = rng.standard_normal(2) * 2
        ref = mvn(mean=[0, 0], cov=[[1, r], [r, 1]]).cdf([h, k])
        assert abs(mv.bvn_cdf(h, k, r) - ref) < 1e-13


def test_fixed_rule_vs_adaptive_and_scipy_genz():
    """2 x 16-point rule vs scipy.quad of the same integrand (its own error) and vs Genz's MVN code at tight
    tolerance (an independent algorithm: mnormt's sadmvn is the same family with abseps 1e-6)."""
    rng = np.random.default_rng(1)
    for it in range(10):
        q = 3 + it % 2
        R = _corr(rng, q, [0.5, 2.0, 5.0][it % 3]); h = rng.standard_normal(q)
        f = mv.pmnorm(h, varcov=R); a = mv.pmnorm(h, varcov=R, adaptive=True)
        s = mvn(mean=np.zeros(q), cov=R, abseps=1e-10, releps=1e-10, maxpts=5_000_000, seed=it).cdf(h)
        assert abs(f - a) < 1e-9
        assert abs(a - s) < 5e-7      # scipy's randomised QMC error, not ours


def test_qei_exact_vs_fast_and_monte_carlo():
    km = to_oracle(small_model("matern5_2", "const", n=30, d=3, seed=13, noise_var=0.05))
    rng = np.random.default_rng(5)
    plugin = float(np.median(km.y))
    for q in (2, 3, 4):
        x = rng.random((q, 3))
        ex = do.qEI(x, km, plugin=plugin, fastCompute=False)
        fa = do.qEI(x, km, plugin=plugin, fastCompute=True)
        assert abs(ex - fa) < 1e-4 * abs(ex) + 1e-9       # the reference's FD form approximates the exact term
        mc, se = do.qEI_monte_carlo(x, km, plugin=plugin, nsim=300000, seed=q)
        assert abs(ex - mc) < 5 * se + 1e-9
        assert ex > 1e-7
    # q = 2 closed form cross-check: qEI of two far-apart independent-ish points is below the sum of EIs
    x = rng.random((2, 3))
    assert do.qEI(x, km, plugin=plugin, fastCompute=False) <= do.EI(x, km, plugin=plugin).sum() + 1e-12


def test_qei_grad_oracle_vs_finite_differences():
    """The reference's own check (R/qEI.grad.R:232-241): analytic gradient vs finite differences of qEI."""
    km = to_oracle(small_model("matern5_2", "linear", n=25, d=2, seed=15, noise_var=0.02))
    rng = np.random.default_rng(1)
    plugin = float(np.median(km.y))
    for q in (2, 3):
        x = rng.random((q, 2))
        g = do.qEI_grad(x, km, plugin=plugin, fastCompute=False)
        h = 1e-5
        fd = np.zeros_like(g)
        for l in range(q):
            for j in range(2):
                xp = x.copy(); xp[l, j] += h; xm = x.copy(); xm[l, j] -= h
                fd[l, j] = (do.qEI(xp, km, plugin=plugin, fastCompute=False) - do.qEI(xm, km, plugin=plugin, fastCompute=False)) / (2 * h)
        assert np.max(np.abs(g - fd)) <= 1e-6 * np.max(np.abs(g))
        gf = do.qEI_grad(x, km, plugin=plugin, fastCompute=True)
        assert np.max(np.abs(g - gf)) <= 1e-6 * np.max(np.abs(g))


def test_lattice_rule_against_low_dimensional_rules_and_monte_carlo():
    """The fixed-lattice Genz rule that serves dimensions 5..10: (i) on 4-variate problems against the deterministic
    Plackett rule (its error class: mean ~2e-5, worst case below 1e-3 on ill-conditioned correlation matrices, which is
    sadmvn's class at the reference's maxpts = 200 q); (ii) on 6- and 9-variate problems against plain Monte-Carlo."""
    rng = np.random.default_rng(0)
    errs = []
    for _ in range(30):
        A = rng.standard_normal((4, 5)); S = A @ A.T
        h = rng.standard_normal(4) * 1.5 * np.sqrt(np.diag(S))
        sd = np.sqrt(np.diag(S))
        errs.append(abs(mv.pmnorm_lattice(h / sd, S / np.outer(sd, sd)) - mv.pmnorm(h, varcov=S)))
    assert max(errs) < 1e-3 and np.mean(errs) < 5e-5
    for q in (6, 9):
        A = rng.standard_normal((q, q + 2)); S = A @ A.T
        sd = np.sqrt(np.diag(S)); h = (rng.standard_normal(q) * 0.5 + 1.0) * sd
        v = mv.pmnorm(h, varcov=S)
        Z = rng.standard_normal((1_000_000, q)) @ np.linalg.cholesky(S).T
        mc = np.mean(np.all(Z <= h, axis=1)); se = np.sqrt(mc * (1 - mc) / 1e6)
        assert abs(v - mc) < 4 * se + 1e-4, (q, v, mc, se)


def test_qei_q6_exact_vs_monte_carlo():
    km = to_oracle(small_model("matern5_2", "const", n=30, d=3, seed=13, noise_var=0.05))
    x = np.random.default_rng(8).random((6, 3))
    plugin = float(np.median(km.y))
    ex = do.qEI(x, km, plugin=plugin, fastCompute=False)
    mc, se = do.qEI_monte_carlo(x, km, plugin=plugin, nsim=300000, seed=6)
    assert abs(ex - mc) < 4 * se + 1e-4 * abs(ex), (ex, mc, se)
