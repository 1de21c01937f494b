"""The oracle's restatement of predict.km (DiceKriging is not under /root/reference and cannot run here) against an
independent implementation of the same posterior: scikit-learn's GaussianProcessRegressor with fixed
hyper-parameters.  DiceKriging's gauss kernel is the anisotropic RBF exp(-sum h_k^2 / (2 theta_k^2)); its
matern5_2 / matern3_2 are tensor products of 1-d Matern kernels, equal to scikit-learn's Matern(nu = 2.5 / 1.5) for
d = 1.  Simple kriging with the trend coefficient as known mean is the zero-mean GP on y - beta.  1e-8 relative
(two different Cholesky routes).  CPU only."""
import numpy as np
import pytest

from oracle import dicekriging as dk

sk = pytest.importorskip("sklearn.gaussian_process")
from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Matern  # noqa: E402


def _check(model, kernel, Xnew, alpha):
    gp = sk.GaussianProcessRegressor(kernel=kernel, alpha=alpha, optimizer=None, normalize_y=False)
    beta = float(model.trend_coef[0])
    gp.fit(model.X, model.y - beta)
    mean, sd = gp.predict(Xnew, return_std=True)
    got = dk.predict_km(model, Xnew, "SK")
    scale = np.max(np.abs(model.y)) + np.sqrt(model.covariance.sd2)
    assert np.max(np.abs(got["mean"] - (mean + beta))) < 1e-8 * scale
    assert np.max(np.abs(got["sd"] ** 2 - sd ** 2)) < 1e-8 * model.covariance.sd2


def test_gauss_kernel_any_dimension():
    rng = np.random.default_rng(0)
    X = rng.random((40, 3)); y = np.sin(3 * X.sum(1)) + X[:, 0]
    theta = np.array([0.4, 0.7, 0.55]); sd2 = 1.7
    m = dk.make_km(X, y, "gauss", theta, sd2, noise_var=np.full(40, 1e-6))
    _check(m, ConstantKernel(sd2, "fixed") * RBF(theta, "fixed"), rng.random((25, 3)), 1e-6)


@pytest.mark.parametrize("name,nu", [("matern5_2", 2.5), ("matern3_2", 1.5), ("exp", 0.5)])
def test_matern_kernels_one_dimension(name, nu):
    rng = np.random.default_rng(1)
    X = np.sort(rng.random((30, 1)), axis=0); y = np.cos(5 * X[:, 0])
    theta, sd2 = 0.3, 0.9
    m = dk.make_km(X, y, name, np.array([theta]), sd2, noise_var=np.full(30, 1e-6))
    _check(m, ConstantKernel(sd2, "fixed") * Matern(theta, "fixed", nu=nu), rng.random((20, 1)), 1e-6)


def test_universal_kriging_is_the_flat_prior_limit():
    """UK with a constant trend = zero-mean GP whose kernel carries an extra constant of huge variance (flat prior on
    beta): the GLS coefficient and the variance inflation term of predict.km(type = "UK") come out of that limit.  1e-5
    (the 1e6 prior variance costs digits)."""
    rng = np.random.default_rng(2)
    X = rng.random((35, 2)); y = 2.0 + np.sin(4 * X[:, 0]) * X[:, 1]
    theta = np.array([0.5, 0.6]); sd2 = 0.8
    m = dk.make_km(X, y, "gauss", theta, sd2, noise_var=np.full(35, 1e-4))      # beta = GLS estimate
    kern = ConstantKernel(sd2, "fixed") * RBF(theta, "fixed") + ConstantKernel(1e6, "fixed")
    gp = sk.GaussianProcessRegressor(kernel=kern, alpha=1e-4, optimizer=None, normalize_y=False).fit(X, y)
    Xn = rng.random((20, 2))
    mean, sd = gp.predict(Xn, return_std=True)
    got = dk.predict_km(m, Xn, "UK")
    assert np.max(np.abs(got["mean"] - mean)) < 1e-5 * (np.max(np.abs(y)) + 1.0)
    assert np.max(np.abs(got["sd"] ** 2 - sd ** 2)) < 1e-5 * sd2
