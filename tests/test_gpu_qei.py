"""GPU parity of qEI (R/qEI.R) and of the device normal CDFs against the oracle, which runs the SAME
deterministic rules (oracle/mvnorm.py).  Tolerances: CDFs 1e-12 absolute; qEI 1e-10 relative with an
absolute floor of 1e-12 x the response scale (the two terms of the formula cancel for poor batches)."""
import ctypes as C

import numpy as np
import pytest

from conftest import small_model, to_oracle

pytestmark = pytest.mark.gpu


def _rand_cov(rng, q):
    A = rng.standard_normal((q, q + 1)); A[:, 0] *= rng.choice([0.3, 2.0, 6.0])
    S = A @ A.T
    return S * rng.uniform(0.2, 3.0)


@pytest.mark.parametrize("q", [1, 2, 3, 4])
def test_device_mvn_cdf_matches_oracle_rules(gpu_lib, q):
    from oracle import mvnorm as mv
    rng = np.random.default_rng(40 + q)
    B = 150
    S = np.stack([_rand_cov(rng, q) for _ in range(B)])
    h = rng.standard_normal((B, q)) * 1.5 * np.sqrt(np.einsum("bii->bi", S))
    out = np.empty(B)
    rc = gpu_lib.dk_mvn_cdf(q, B, C.c_void_p(np.ascontiguousarray(h).ctypes.data),
                            C.c_void_p(np.ascontiguousarray(S).ctypes.data), C.c_void_p(out.ctypes.data))
    assert rc == 0, gpu_lib.dk_last_error()
    ref = np.array([mv.pmnorm(h[b], varcov=S[b]) for b in range(B)])
    assert np.max(np.abs(out - ref)) < 1e-12


def _check_qei(km, xb, **kw):
    from diceoptim_b200 import qEI
    from oracle import diceoptim as do
    ko = to_oracle(km)
    got = qEI(xb, km, **kw)
    ref = np.array([do.qEI(xb[b], ko, fastCompute=False, **kw) for b in range(xb.shape[0])])
    scale = np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2)
    assert np.all(np.abs(got - ref) <= 1e-10 * np.abs(ref) + 1e-12 * scale), np.max(np.abs(got - ref))
    return got, ref


@pytest.mark.parametrize("q", [2, 3, 4])
@pytest.mark.parametrize("trend,type", [("const", "UK"), ("const", "SK"), ("linear", "UK")])
def test_qei_matches_oracle(gpu_lib, q, trend, type):
    km = small_model("matern5_2", trend, n=140, d=3, seed=50 + q)
    rng = np.random.default_rng(q)
    xb = rng.random((40, q, 3))
    plugin = float(np.median(km.y))                 # a reachable target: EI terms of order 0.1
    got, ref = _check_qei(km, xb, plugin=plugin, type=type)
    assert np.max(ref) > 1e-3 and np.all(ref >= 0)
    _check_qei(km, xb[:10], type=type)             # default plugin = min(y)
    _check_qei(km, xb[:10], plugin=plugin, type=type, minimization=False)
    km.release()


def test_qei_single_batch_and_other_kernels(gpu_lib):
    from diceoptim_b200 import qEI
    from oracle import diceoptim as do
    for name in ("gauss", "exp", "matern3_2", "powexp"):
        km = small_model(name, "const", n=60, d=2, seed=3, noise_var=0.01 if name == "gauss" else None)
        x = np.random.default_rng(1).random((3, 2))
        got = qEI(x, km, plugin=float(np.median(km.y)))
        assert isinstance(got, float)
        ref = do.qEI(x, to_oracle(km), plugin=float(np.median(km.y)), fastCompute=False)
        assert abs(got - ref) <= 1e-10 * abs(ref) + 1e-12
        km.release()


def test_qei_duplicates_follow_reference_unique_rule(gpu_lib):
    """R/qEI.R:124-134: points equal (to 8 digits) to a design row or to an earlier batch point are dropped;
    nothing left -> 0; one left -> EI."""
    from diceoptim_b200 import qEI
    from oracle import diceoptim as do
    km = small_model("matern5_2", "const", n=50, d=2, seed=9)
    ko = to_oracle(km)
    rng = np.random.default_rng(4)
    plugin = float(np.median(km.y))
    fresh = rng.random((4, 2))
    cases = [
        np.vstack([km.X[3], fresh[0], fresh[1], fresh[2]]),            # one design row  -> q' = 3
        np.vstack([fresh[0], fresh[0] + 1e-10, fresh[1], fresh[2]]),   # in-batch duplicate after rounding
        np.vstack([km.X[0], km.X[1], km.X[2], fresh[3]]),              # q' = 1 -> EI
        np.vstack([km.X[0], km.X[1], km.X[2], km.X[3]]),               # all duplicates -> 0
    ]
    xb = np.stack(cases)
    got = qEI(xb, km, plugin=plugin)
    ref = np.array([do.qEI(c, ko, plugin=plugin, fastCompute=False) for c in cases])
    assert got[3] == 0.0 and ref[3] == 0.0
    assert np.all(np.abs(got - ref) <= 1e-10 * np.abs(ref) + 1e-12)
    with pytest.raises(Exception):
        qEI(rng.random((11, 2)), km)                                     # q = 11 > DK_MAX_Q_LATTICE: refused loudly
    km.release()


def test_qei_against_monte_carlo(gpu_lib):
    """The reference's own cross-check (R/qEI.R:89-104): qEI vs the mean improvement over draws."""
    from diceoptim_b200 import qEI
    from oracle import diceoptim as do
    km = small_model("matern5_2", "const", n=80, d=3, seed=13)
    x = np.random.default_rng(2).random((4, 3))
    plugin = float(np.median(km.y))
    got = qEI(x, km, plugin=plugin)
    mc, se = do.qEI_monte_carlo(x, to_oracle(km), plugin=plugin, nsim=400000)
    assert abs(got - mc) < 5 * se + 1e-6
    km.release()


@pytest.mark.parametrize("q", [2, 3, 4])
@pytest.mark.parametrize("trend,type", [("const", "UK"), ("const", "SK"), ("linear", "UK")])
def test_qei_grad_matches_oracle(gpu_lib, q, trend, type):
    """qEI.grad (R/qEI.grad.R) vs the oracle's line-by-line restatement (itself checked against finite
    differences of qEI in tests/test_oracle_mvnorm.py); 1e-10 of the largest gradient entry."""
    from diceoptim_b200 import qEI_grad
    from oracle import diceoptim as do
    km = small_model("matern5_2", trend, n=70, d=3, seed=60 + q, noise_var=0.02)
    ko = to_oracle(km)
    rng = np.random.default_rng(10 + q)
    xb = rng.random((12, q, 3))
    plugin = float(np.median(km.y))
    got = qEI_grad(xb, km, plugin=plugin, type=type)
    assert got.shape == (12, q, 3)
    for b in range(12):
        ref = do.qEI_grad(xb[b], ko, plugin=plugin, type=type, fastCompute=False)
        assert np.max(np.abs(got[b] - ref)) <= 1e-10 * np.max(np.abs(ref)) + 1e-13, (b, got[b], ref)
    gm = qEI_grad(xb[:3], km, plugin=-plugin, type=type, minimization=False)
    for b in range(3):
        ref = do.qEI_grad(xb[b], ko, plugin=-plugin, type=type, minimization=False, fastCompute=False)
        assert np.max(np.abs(gm[b] - ref)) <= 1e-10 * np.max(np.abs(ref)) + 1e-13
    km.release()


@pytest.mark.parametrize("q", [3, 5])
def test_qei_envir_reuse_gives_the_gradient_of_the_same_pass(gpu_lib, q):
    """qEI(x, envir) then qEI.grad(x, envir) (R/qEI.R:190-198, R/qEI.grad.R:299-316): one pass over the model; the gradient
    equals the separate call's bit for bit, the value is qEI at x itself (the separate call rounds x to 8 digits first)."""
    from diceoptim_b200 import qEI, qEI_grad
    km = small_model("matern5_2", "const", n=60, d=2, seed=17)
    xb = np.random.default_rng(3).random((4, q, 2))
    pl = float(np.median(km.y))
    env = {}
    v = qEI(xb, km, plugin=pl, envir=env)
    g = qEI_grad(xb, km, plugin=pl, envir=env)
    assert g is env["qEI.grad"]
    assert np.array_equal(g, qEI_grad(xb, km, plugin=pl))
    assert np.allclose(v, qEI(xb, km, plugin=pl), rtol=0, atol=1e-7)           # x moved by up to 5e-9 in the separate call
    assert np.array_equal(qEI(xb, km, plugin=pl), qEI(np.round(xb, 8), km, plugin=pl, envir={}))   # same x -> same bits
    g2 = qEI_grad(xb + 1e-3, km, plugin=pl, envir=env)                             # another x: computed, not reused
    assert g2 is not env["qEI.grad"] and not np.array_equal(g2, g)
    km.release()


def test_qei_grad_guards(gpu_lib):
    """Zero matrix when a batch point duplicates the design (R/qEI.grad.R:293-295); q = 1 -> EI.grad."""
    from diceoptim_b200 import qEI_grad, EI_grad
    km = small_model("matern5_2", "const", n=40, d=2, seed=3)
    rng = np.random.default_rng(0)
    x = np.vstack([km.X[5], rng.random((2, 2))])
    assert np.all(qEI_grad(x, km, plugin=float(np.median(km.y))) == 0.0)
    x1 = rng.random((1, 2))
    assert np.allclose(qEI_grad(x1, km, plugin=0.1), EI_grad(x1[0], km, plugin=0.1)[None, :], rtol=0, atol=0)
    km.release()


# ---- q = 5 .. 10: the fixed-lattice Genz rule (csrc/dk_qei_gen.cuh, oracle.mvnorm.pmnorm_lattice) ----
@pytest.mark.parametrize("q", [5, 7, 10])
def test_device_lattice_cdf_matches_oracle_rule(gpu_lib, q):
    """same rule on both sides (reordering, Cholesky, 8192 Richtmyer points): rounding-level agreement"""
    from oracle import mvnorm as mv
    rng = np.random.default_rng(60 + q)
    B = 12
    S = np.stack([_rand_cov(rng, q) for _ in range(B)])
    h = (rng.standard_normal((B, q)) * 0.8 + 0.6) * np.sqrt(np.einsum("bii->bi", S))
    out = np.empty(B)
    rc = gpu_lib.dk_mvn_cdf(q, B, C.c_void_p(np.ascontiguousarray(h).ctypes.data),
                            C.c_void_p(np.ascontiguousarray(S).ctypes.data), C.c_void_p(out.ctypes.data))
    assert rc == 0, gpu_lib.dk_last_error()
    ref = np.array([mv.pmnorm(h[b], varcov=S[b]) for b in range(B)])
    assert np.max(np.abs(out - ref)) < 1e-11, np.max(np.abs(out - ref))
    assert ref.max() > 1e-3


@pytest.mark.parametrize("q,trend,type", [(5, "const", "UK"), (6, "linear", "UK"), (8, "const", "SK"), (10, "const", "UK")])
def test_qei_general_q_matches_oracle(gpu_lib, q, trend, type):
    km = small_model("matern5_2", trend, n=140, d=3, seed=70 + q)
    ko = to_oracle(km)
    from diceoptim_b200 import qEI
    from oracle import diceoptim as do
    rng = np.random.default_rng(q)
    B = 9
    xb = rng.random((B, q, 3))
    xb[1, 2] = km.X[5]                       # a design row: the batch shrinks to q - 1 points
    xb[2, 3] = xb[2, 0] + 1e-10              # an in-batch duplicate after rounding
    xb[3, 1:] = km.X[1:q]                    # q' = 1 -> EI
    plugin = float(np.median(km.y))
    got = qEI(xb, km, plugin=plugin, type=type)
    ref = np.array([do.qEI(xb[b], ko, plugin=plugin, type=type, fastCompute=False) for b in range(B)])
    scale = np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2)
    assert np.all(np.abs(got - ref) <= 1e-9 * np.abs(ref) + 1e-11 * scale), np.max(np.abs(got - ref))
    assert np.max(ref) > 1e-3
    gm = qEI(xb[:4], km, plugin=plugin, type=type, minimization=False)
    rm = np.array([do.qEI(xb[b], ko, plugin=plugin, type=type, minimization=False, fastCompute=False) for b in range(4)])
    assert np.all(np.abs(gm - rm) <= 1e-9 * np.abs(rm) + 1e-11 * scale)
    km.release()


@pytest.mark.parametrize("q,trend,type", [(5, "const", "UK"), (5, "linear", "UK"), (6, "const", "SK"), (7, "const", "UK")])
def test_qei_grad_general_q_matches_oracle(gpu_lib, q, trend, type):
    """qEI.grad for q > 4 (csrc/dk_qei_grad_gen.cuh) against the oracle's restatement of R/qEI.grad.R:267-424, which evaluates
    its normal CDFs by the same rules (deterministic up to dimension 4, the fixed lattice above); the q - 3 .. q variate CDFs
    of GPhi / HPhi / term B all take part.  1e-9 of the largest entry (sums of a few hundred CDF terms of both signs; measured
    5e-12 at cfg3's model, profiles/parity_r02.md)."""
    from diceoptim_b200 import qEI_grad
    from oracle import diceoptim as do
    km = small_model("matern5_2", trend, n=60, d=2, seed=80 + q, noise_var=0.02)
    ko = to_oracle(km)
    rng = np.random.default_rng(20 + q)
    B = 3 if q <= 6 else 2
    xb = rng.random((B, q, 2))
    plugin = float(np.median(km.y))
    got = qEI_grad(xb, km, plugin=plugin, type=type)
    assert got.shape == (B, q, 2)
    for b in range(B):
        ref = do.qEI_grad(xb[b], ko, plugin=plugin, type=type, fastCompute=False)
        assert np.max(np.abs(ref)) > 1e-6
        assert np.max(np.abs(got[b] - ref)) <= 1e-9 * np.max(np.abs(ref)), (b, np.max(np.abs(got[b] - ref)), np.max(np.abs(ref)))
    if q == 5:
        gm = qEI_grad(xb[:1], km, plugin=-plugin, type=type, minimization=False)
        ref = do.qEI_grad(xb[0], ko, plugin=-plugin, type=type, minimization=False, fastCompute=False)
        assert np.max(np.abs(gm[0] - ref)) <= 1e-9 * np.max(np.abs(ref))
        xd = xb[:1].copy(); xd[0, 3] = km.X[7]                     # a duplicate of the design -> zero matrix
        assert np.all(qEI_grad(xd, km, plugin=plugin, type=type) == 0.0)
    km.release()


def test_qei_grad_general_q_is_the_derivative_of_the_device_value(gpu_lib):
    """Independent of the oracle: central differences of the device qEI (q = 6; its CDFs come from the same fixed lattice, so
    the value is a smooth function of x) against the device gradient."""
    from diceoptim_b200 import qEI, qEI_grad
    km = small_model("matern5_2", "const", n=50, d=2, seed=91)
    x = np.random.default_rng(5).random((6, 2))
    plugin = float(np.median(km.y))
    g = qEI_grad(x, km, plugin=plugin)
    h = 1e-5
    xs = []
    for l in range(6):
        for j in range(2):
            for sgn in (1.0, -1.0):
                xp = x.copy(); xp[l, j] += sgn * h
                xs.append(xp)
    v = qEI(np.stack(xs), km, plugin=plugin).reshape(6, 2, 2)
    fd = (v[:, :, 0] - v[:, :, 1]) / (2 * h)
    assert np.max(np.abs(g)) > 1e-4
    # the lattice value carries a 1e-5-relative quadrature error whose derivative is not controlled: 1e-3 of the largest entry
    assert np.max(np.abs(fd - g)) <= 1e-3 * np.max(np.abs(g)), (fd, g)
    km.release()


def test_qei_q10_against_monte_carlo_reference_example(gpu_lib):
    """The reference's own self-check (R/qEI.R:66-115): hartman4, 3^4 = 81-point factorial design, q = 10, exact qEI against
    the Monte-Carlo mean improvement over draws of the conditional Gaussian vector."""
    from diceoptim_b200 import qEI, synth
    from oracle import diceoptim as do
    g = np.array([0.0, 0.5, 1.0])
    X = np.array(np.meshgrid(g, g, g, g, indexing="ij")).reshape(4, -1).T
    y = synth.hartman4(X)
    km = synth.make_km(X, y, "matern5_2", np.full(4, 0.6), float(np.var(y)), trend="const")
    x = np.random.default_rng(7).random((10, 4))
    got = qEI(x, km)
    mc, se = do.qEI_monte_carlo(x, to_oracle(km), nsim=600000)
    assert got > 0 and abs(got - mc) < 4 * se + 2e-5 * (np.max(np.abs(y)) + 1), (got, mc, se)
    km.release()
