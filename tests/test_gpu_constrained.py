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


def _noisy_models(n=30, d=2):
    """Few, noisy observations: every term of the SUR criterion (bivariate CDFs included) is of order 1e-2 .. 1."""
    from diceoptim_b200 import synth
    X = synth.lhs(n, d, 11)
    yo = np.sin(4.0 * X[:, 0]) + X[:, 1]
    yc1 = X[:, 0] - 0.5 + 0.3 * np.cos(5 * X[:, 1])
    yc2 = 0.4 - X[:, 1] + 0.2 * X[:, 0]
    yc3 = 0.2 * np.sin(6 * X[:, 0] * X[:, 1])
    nv = np.full(n, 0.05)
    fun = synth.make_km(X, yo, "matern5_2", np.full(d, 0.4), float(np.var(yo)), noise_var=nv)
    c1 = synth.make_km(X, yc1, "matern3_2", np.full(d, 0.5), float(np.var(yc1)), noise_var=nv)
    c2 = synth.make_km(X, yc2, "gauss", np.full(d, 0.5), float(np.var(yc2)), trend="linear", noise_var=nv)
    c3 = synth.make_km(X, yc3, "exp", np.full(d, 0.6), float(np.var(yc3)), nugget=0.01)
    return fun, [c1, c2, c3]


@pytest.mark.parametrize("ncst,eq", [(1, False), (2, False), (3, False), (1, [True]), (2, [False, True]), (3, [True, False, False])])
@pytest.mark.parametrize("type", ["UK", "SK"])
def test_crit_sur_cst_matches_oracle(gpu_lib, ncst, eq, type):
    """crit_SUR_cst over a batch vs the oracle's restatement of R/crit_SUR_cst.R row by row; 1e-9 relative with an
    absolute floor of 1e-13 (the criterion is a difference of probabilities times 1 / M^2)."""
    from diceoptim_b200 import crit_SUR_cst
    from oracle import diceoptim as do
    fun, cst = _noisy_models()
    cst = cst[:ncst]
    rng = np.random.default_rng(3)
    P = rng.random((70, 2))
    x = rng.random((24, 2))
    x[0] = fun.X[2]                                # on the design (noisy models: not refused, sk small)
    cc = {"integration.points": P}
    got = crit_SUR_cst(x, fun, cst, equality=eq, critcontrol=cc, type=type)
    fo, co = to_oracle(fun), [to_oracle(m) for m in cst]
    ref = np.array([do.crit_SUR_cst(xi, fo, co, equality=eq, critcontrol=cc, type=type) for xi in x])
    assert np.max(np.abs(ref)) > 1e-7              # the case exercises the formula, not rounding noise
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-13), np.max(np.abs(got - ref))
    for m in cst + [fun]:
        m.release()


def test_crit_sur_cst_weights_no_feasible_and_guards(gpu_lib):
    from diceoptim_b200 import crit_SUR_cst
    from oracle import diceoptim as do
    fun, cst = _noisy_models()
    rng = np.random.default_rng(8)
    P = rng.random((50, 2)); w = rng.random(50); w /= w.sum()
    x = rng.random((10, 2))
    fo, co = to_oracle(fun), [to_oracle(m) for m in cst[:2]]
    cc = {"integration.points": P, "integration.weights": w, "tolConstraints": [0.02, 0.0]}
    got = crit_SUR_cst(x, fun, cst[:2], critcontrol=cc)
    ref = np.array([do.crit_SUR_cst(xi, fo, co, critcontrol=cc) for xi in x])
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-13)
    # no feasible observation: shift a constraint's observations up so that obj.min = Inf (:184)
    from diceoptim_b200 import synth
    bad = synth.make_km(cst[0].X, cst[0].y + 10.0, "matern3_2", np.full(2, 0.5), float(np.var(cst[0].y)), noise_var=np.full(30, 0.05))
    got = crit_SUR_cst(x, fun, [bad], critcontrol={"integration.points": P})
    ref = np.array([do.crit_SUR_cst(xi, fo, [to_oracle(bad)], critcontrol={"integration.points": P}) for xi in x])
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-13)
    assert isinstance(crit_SUR_cst(x[0], fun, cst[:1], critcontrol={"integration.points": P}), float)
    with pytest.raises(ValueError):
        crit_SUR_cst(x, fun, cst[:1])
    # integration points drawn by integration_design_cst (R/crit_SUR_cst.R:139-143), plain and importance-sampled
    from diceoptim_b200 import integration_design_cst
    v = crit_SUR_cst(x, fun, cst[:1], critcontrol={"lower": [0, 0], "upper": [1, 1]})
    assert v.shape == (10,) and np.all(np.isfinite(v))
    des = integration_design_cst({"n.points": 64, "distrib": "SUR", "n.candidates": 512}, [0, 0], [1, 1], fun, cst[:2], seed=1)
    assert des["integration.points"].shape == (64, 2) and des["integration.weights"].shape == (64,)
    assert np.all(des["integration.weights"] > 0)
    ci = {"integration.points": des["integration.points"], "integration.weights": des["integration.weights"]}
    got = crit_SUR_cst(x, fun, cst[:2], critcontrol=ci)
    ref = np.array([do.crit_SUR_cst(xi, fo, co, critcontrol=ci) for xi in x])
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-13)
    assert integration_design_cst({"n.points": 32, "distrib": "sobol"}, [0, 0], [2, 1])["integration.points"].max() <= 2.0
    with pytest.warns(UserWarning):
        assert np.all(np.isnan(crit_SUR_cst(x, fun, cst + [cst[0]], critcontrol={"integration.points": P})))
    for m in cst + [fun, bad]:
        m.release()


def test_constrained_batch_equals_row_by_row(gpu_lib):
    """Position independence across ragged tiles: a row evaluated inside a batch of 131 (tiles of 64 + 64 + 3) and on
    its own gives the same bits, for the distances, crit_EFI and crit_SUR_cst (129 integration points: ragged too)."""
    from diceoptim_b200 import checkPredict, crit_EFI, crit_SUR_cst
    fun, cst = _noisy_models()
    cst = cst[:2]
    rng = np.random.default_rng(21)
    x = rng.random((131, 2)); P = rng.random((129, 2))
    cc = {"integration.points": P}
    full_d = checkPredict(x, cst + [fun], return_distance=True)
    full_e = crit_EFI(x, fun, cst)
    full_s = crit_SUR_cst(x, fun, cst, critcontrol=cc)
    for i in (0, 63, 64, 127, 128, 130):
        assert checkPredict(x[i], cst + [fun], return_distance=True)[0] == full_d[i]
        assert crit_EFI(x[i], fun, cst) == full_e[i]
        assert crit_SUR_cst(x[i], fun, cst, critcontrol=cc) == full_s[i]
    again = crit_SUR_cst(x, fun, cst, critcontrol=cc)
    assert np.array_equal(again, full_s)
    for m in cst + [fun]:
        m.release()


def test_constrained_edge_inputs(gpu_lib):
    """Empty candidate sets return empty vectors; a NaN row leaves its tile neighbours alone (and is NaN itself for crit_EFI)."""
    from diceoptim_b200 import checkPredict, crit_EFI, crit_SUR_cst, prob_of_feasible_improvement
    fun, cst = _noisy_models()
    cst = cst[:2]
    P = np.random.default_rng(2).random((40, 2))
    cc = {"integration.points": P}
    empty = np.empty((0, 2))
    assert checkPredict(empty, cst + [fun]).shape == (0,)
    assert crit_EFI(empty, fun, cst).shape == (0,)
    assert prob_of_feasible_improvement(empty, fun, cst).shape == (0,)
    assert crit_SUR_cst(empty, fun, cst, critcontrol=cc).shape == (0,)
    x = np.random.default_rng(4).random((70, 2))
    clean_e = crit_EFI(x, fun, cst); clean_s = crit_SUR_cst(x, fun, cst, critcontrol=cc)
    xb = x.copy(); xb[5, 1] = np.nan
    e = crit_EFI(xb, fun, cst); s = crit_SUR_cst(xb, fun, cst, critcontrol=cc)
    keep = np.arange(70) != 5
    assert np.isnan(e[5]) and np.array_equal(e[keep], clean_e[keep])
    assert np.array_equal(s[keep], clean_s[keep])     # the reference stops on a NaN point (if (NA > -10)); row 5 is unspecified
    for m in cst + [fun]:
        m.release()
