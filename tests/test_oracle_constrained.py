"""CPU checks of the constrained-criteria restatements: test_feas_vec cases, and the identity the engine uses for
checkPredict's covdist, kn_xy = kxy - kxy K^-1 kyy = delta .* T^-1 T'^-1 kxy (delta = noise.var or 0)."""
import numpy as np
from scipy.linalg import solve_triangular

from oracle import dicekriging as dk
from oracle import diceoptim as do


def _km(noise):
    rng = np.random.default_rng(0)
    X = rng.random((30, 2)); y = np.sin(4 * X[:, 0]) + X[:, 1]
    kw = dict(noise_var=np.linspace(0.01, 0.05, 30)) if noise else dict(nugget=1e-3)
    return dk.make_km(X, y, "matern5_2", np.array([0.4, 0.5]), float(np.var(y)), **kw)


def test_feas_vec_cases():
    cst = np.array([[-1.0, -0.01], [0.2, -1.0], [-0.5, 0.04], [-0.5, -0.2]])
    assert do.test_feas_vec(cst).tolist() == [True, False, False, True]
    assert do.test_feas_vec(cst, equality=[False, True]).tolist() == [True, False, True, False]
    assert do.test_feas_vec(cst, equality=[False, True], tolConstraints=[0.3, 0.01]).tolist() == [True, False, False, False]
    assert do.test_feas_vec(np.array([-1.0, -2.0])).tolist() == [True]


def test_covdist_identity():
    for noise in (True, False):
        m = _km(noise)
        x = np.random.default_rng(1).random((7, 2))
        kxy = dk.cov_mat1_mat2(m.covariance, x, m.X)
        kyy = dk.cov_matrix(m.covariance, m.X)
        kn_xy = kxy - kxy @ np.linalg.inv(m.T.T @ m.T) @ kyy
        v = solve_triangular(m.T, kxy.T, trans="T", lower=False)
        w = solve_triangular(m.T, v, lower=False)
        delta = m.noise_var if noise else np.zeros(m.X.shape[0])
        assert np.max(np.abs(kn_xy - (delta[:, None] * w).T)) < 1e-10 * m.covariance.sd2
        d = do.check_predict_distance(x, [m], "covdist")
        assert np.all(np.isfinite(d)) and np.all(d >= 0)


def test_crit_sur_cst_against_its_definition_by_monte_carlo():
    """One inequality constraint: the criterion is the expected drop, after observing f and g at x, of
    mean_u w_u P(g(u) <= 0) P(f(u) <= best feasible value), where the best feasible value becomes min(obj.min, f(x))
    when g(x) <= 0.  Here the expectation is a 3e5-draw Monte-Carlo over (f(x), g(x)) with the Gaussian conditioning
    formulas and predict.km's own posterior covariances, instead of the pbivnorm algebra of R/crit_SUR_cst.R."""
    from scipy.special import ndtr
    rng = np.random.default_rng(5)
    X = rng.random((18, 2))
    yo = np.sin(4 * X[:, 0]) + X[:, 1] + 0.05 * rng.standard_normal(18)
    yc = X[:, 0] - 0.45 + 0.3 * np.cos(5 * X[:, 1]) + 0.05 * rng.standard_normal(18)
    nv = np.full(18, 0.03)
    fun = dk.make_km(X, yo, "matern5_2", np.array([0.4, 0.5]), float(np.var(yo)), noise_var=nv)
    cst = dk.make_km(X, yc, "matern3_2", np.array([0.5, 0.5]), float(np.var(yc)), noise_var=nv)
    P = rng.random((25, 2)); Mi = P.shape[0]
    feas = do.test_feas_vec(yc[:, None])
    obj_min = float(np.min(yo[feas]))
    S = 300_000
    informative = 0
    for x in rng.random((2, 2)):
        got = do.crit_SUR_cst(x, fun, [cst], critcontrol={"integration.points": P, "distance": "covratio", "threshold": 1e-9})
        tot = np.zeros(S)
        parts = {}
        for tag, m in (("f", fun), ("g", cst)):
            pr = dk.predict_km(m, np.vstack([P, x]), "UK", cov_compute=True)
            mn, sn2 = pr["mean"][:-1], np.diag(pr["cov"])[:-1]
            mk, sk2, kn = pr["mean"][-1], pr["cov"][-1, -1], pr["cov"][:-1, -1]
            ynew = mk + np.sqrt(sk2) * rng.standard_normal(S)
            parts[tag] = (mn, sn2, mk, sk2, kn, ynew)
        mn_f, sn2_f, mk_f, sk2_f, kn_f, yf = parts["f"]
        mn_g, sn2_g, mk_g, sk2_g, kn_g, yg = parts["g"]
        T = np.where(yg <= 0.0, np.minimum(obj_min, yf), obj_min)
        for u in range(Mi):
            old = ndtr((obj_min - mn_f[u]) / np.sqrt(sn2_f[u])) * ndtr((0.0 - mn_g[u]) / np.sqrt(sn2_g[u]))
            mf = mn_f[u] + kn_f[u] / sk2_f * (yf - mk_f); sf = np.sqrt(max(sn2_f[u] - kn_f[u] ** 2 / sk2_f, 1e-300))
            mg = mn_g[u] + kn_g[u] / sk2_g * (yg - mk_g); sg = np.sqrt(max(sn2_g[u] - kn_g[u] ** 2 / sk2_g, 1e-300))
            new = ndtr((T - mf) / sf) * ndtr((0.0 - mg) / sg)
            tot += (old - new) * (1.0 / Mi) / Mi          # mean(a * weights), weights = 1 / Mi
        mc, se = tot.mean(), tot.std() / np.sqrt(S)
        assert abs(got - mc) < 5 * se + 1e-12, (got, mc, se)
        informative += got > 20 * se
    assert informative >= 1          # at least one point where the criterion is resolved to a few percent
