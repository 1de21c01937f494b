"""The batched projected L-BFGS (diceoptim_b200/optim.py) against scipy's L-BFGS-B, one problem at a time, on
box-constrained quadratics (active bounds at the optimum) and on the Rosenbrock valley.  CPU only."""
import numpy as np
from scipy.optimize import minimize

from diceoptim_b200.optim import batched_lbfgsb


def test_box_quadratics_match_scipy():
    rng = np.random.default_rng(0)
    S, D = 40, 6
    A = rng.standard_normal((S, D, D)); A = np.einsum("sij,skj->sik", A, A) + 0.5 * np.eye(D)
    b = rng.standard_normal((S, D)) * 3.0
    lo, up = np.full(D, -0.5), np.full(D, 0.7)

    def fg(X):
        Ax = np.einsum("sij,sj->si", A, X)
        return 0.5 * np.einsum("si,si->s", X, Ax) - np.einsum("si,si->s", b, X), Ax - b

    x0 = rng.uniform(lo, up, size=(S, D))
    res = batched_lbfgsb(fg, x0, lo, up, maxit=200, factr=10.0)
    assert np.all(res["x"] >= lo) and np.all(res["x"] <= up)
    for s in range(S):
        ref = minimize(lambda t: 0.5 * t @ A[s] @ t - b[s] @ t, x0[s], jac=lambda t: A[s] @ t - b[s], method="L-BFGS-B",
                       bounds=list(zip(lo, up)), options={"ftol": 1e-15, "gtol": 1e-10, "maxiter": 500})
        assert res["f"][s] <= ref.fun + 1e-8 * max(1.0, abs(ref.fun)), (s, res["f"][s], ref.fun)
        assert np.max(np.abs(res["x"][s] - ref.x)) < 1e-4


def test_rosenbrock_starts_reach_the_minimum():
    def fg(X):
        a, b = X[:, :-1], X[:, 1:]
        f = np.sum(100.0 * (b - a * a) ** 2 + (1.0 - a) ** 2, axis=1)
        g = np.zeros_like(X)
        g[:, :-1] += -400.0 * a * (b - a * a) - 2.0 * (1.0 - a)
        g[:, 1:] += 200.0 * (b - a * a)
        return f, g

    rng = np.random.default_rng(1)
    x0 = rng.uniform(-1.5, 1.5, size=(16, 4))
    res = batched_lbfgsb(fg, x0, -2.0, 2.0, maxit=500, factr=10.0)
    # the 4-d Rosenbrock function has a second local minimum near (-0.78, 0.61, 0.37, 0.14), f = 3.70
    assert np.all((res["f"] < 1e-8) | (np.abs(res["f"] - 3.7014) < 1e-3)), res["f"]
    assert np.sum(res["f"] < 1e-8) >= 8
    assert res["evaluations"] < 2000          # launches, not launches x starts


def test_converged_rows_stay_put_and_nan_rows_do_not_poison():
    def fg(X):
        f = np.sum(X * X, axis=1); g = 2.0 * X
        f = np.where(X[:, 0] > 10.0, np.nan, f)
        return f, g

    x0 = np.array([[0.0, 0.0], [1.0, -1.0], [0.5, 0.25]])
    res = batched_lbfgsb(fg, x0, -2.0, 2.0, maxit=50)
    assert np.allclose(res["x"], 0.0, atol=1e-6) and res["iterations"][0] == 0
