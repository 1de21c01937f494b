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
