"""Constrained criteria on the GPU (checkPredict, crit_EFI, prob.of.feasible.improvement over several km models)
against the oracle's restatements of R/checkPredict.R, R/crit_EFI.R, R/prob.of.feasible.improvement.R.
Tolerances: distances 1e-9 relative with an absolute floor of 1e-6 (mindist is the square root of a difference of
quantities of order sd2 that cancel to ~1e-13 sd2 next to a design point of a noise-free model, where the oracle's
rounding noise in kn_xy decides between 0 and 1e-7); products 1e-10 relative."""
import numpy as np
import pytest

from conftest import small_model, to_oracle

pytestmark = pytest.mark.gpu


def _models(noise=True, n=50, d=3):
    from diceoptim_b200 import synth
    rng = np.random.default_rng(5)
    X = synth.lhs(n, d, 7)
    yo = np.sin(3.0 * X.sum(1)) + 0.5 * X[:, 0] ** 2
    yc1 = X[:, 0] - 0.6 + 0.2 * np.cos(4 * X[:, 1])
    yc2 = 0.3 - X[:, 1] * X[:, 2]
    kw = dict(noise_var=np.full(n, 0.02)) if noise else {}
    fun = synth.make_km(X, yo, "matern5_2", np.full(d, 0.5), float(np.var(yo)), **kw)
    c1 = synth.make_km(X, yc1, "matern3_2", np.full(d, 0.6), float(np.var(yc1)), trend="linear", **kw)
    c2 = synth.make_km(X, yc2, "gauss", np.full(d, 0.4), float(np.var(yc2)), **({"nugget": 1e-3} if not noise else kw))
    return fun, [c1, c2], rng


@pytest.mark.parametrize("distance", ["euclidean", "covdist", "covratio"])
@pytest.mark.parametrize("type", ["UK", "SK"])
@pytest.mark.parametrize("noise", [True, False])
def test_check_predict_distances(gpu_lib, distance, type, noise):
    from diceoptim_b200 import checkPredict
    from oracle import diceoptim as do
    fun, cst, rng = _models(noise)
    x = rng.random((150, 3))
    x[:5] = fun.X[:5]                              # on the design
    x[5:10] = fun.X[5:10] + 1e-7                   # almost on it
    models = cst + [fun]
    got = checkPredict(x, models, distance=distance, type=type, return_distance=True)
    ref = do.check_predict_distance(x, [to_oracle(m) for m in models], distance=distance, type=type)
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-6), np.max(np.abs(got - ref))
    flags = checkPredict(x, models, distance=distance, type=type, threshold=1e-3)
    far = np.abs(ref - 1e-3) > 2e-6               # away from the threshold the verdicts agree exactly
    assert np.array_equal(flags[far], (ref < 1e-3)[far])
    assert not checkPredict(x, models, distance="none").any()
    for m in models:
        m.release()


def test_crit_efi_and_probability(gpu_lib):
    from diceoptim_b200 import crit_EFI, prob_of_feasible_improvement
    from oracle import diceoptim as do
    fun, cst, rng = _models(True)
    fo, co = to_oracle(fun), [to_oracle(m) for m in cst]
    x = rng.random((60, 3))
    x[0] = fun.X[3]
    for eq, cc in [(False, None), ([False, True], None),
                   ([True, False], {"tolConstraints": [0.1, 0.0], "distance": "covratio", "threshold": 1e-3})]:
        got = crit_EFI(x, fun, cst, equality=eq, critcontrol=cc)
        ref = np.array([do.crit_EFI(xi, fo, co, equality=eq, critcontrol=cc) for xi in x])
        assert np.all(np.abs(got - ref) <= 1e-10 * np.abs(ref) + 1e-16), (eq, np.max(np.abs(got - ref)))
        gp = prob_of_feasible_improvement(x, fun, cst, equality=eq, critcontrol=cc)
        rp = do.prob_of_feasible_improvement(x, fo, co, equality=eq, critcontrol=cc)
        assert np.all(np.abs(gp - rp) <= 1e-10 * np.abs(rp) + 1e-16)
    assert isinstance(crit_EFI(x[1], fun, cst), float)
    # no feasible observation: the improvement factor is 1 (R/crit_EFI.R:160-161)
    got = crit_EFI(x, fun, cst, plugin=np.inf)
    ref = np.array([do.crit_EFI(xi, fo, co, plugin=np.inf) for xi in x])
    assert np.all(np.abs(got - ref) <= 1e-10 * np.abs(ref) + 1e-16)
    for m in cst + [fun]:
        m.release()


def test_feas_single_model_and_errors(gpu_lib):
    from diceoptim_b200 import Feas, checkPredict, crit_EFI
    from oracle import diceoptim as do
    km = small_model("matern5_2", "const", n=40, d=2, seed=4)
    x = np.random.default_rng(0).random((30, 2))
    for eq, thr in [(False, 0.0), (False, 0.3), (True, 0.05)]:
        assert np.allclose(Feas(km, x, equality=eq, threshold=thr), do.Feas(to_oracle(km), x, equality=eq, threshold=thr),
                           rtol=1e-10, atol=1e-15)
    with pytest.raises(ValueError):
        checkPredict(x, [km], distance="manhattan")
    with pytest.raises(TypeError):
        crit_EFI(x, km, [object()])
    km.release()
