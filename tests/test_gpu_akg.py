"""GPU parity of the batched AKG (R/AKG.R) against the oracle's line-by-line restatement.  The engine uses
the exact identity cn = noise.var .* T^-1 (v + M r) where the reference subtracts two nearly equal vectors,
so agreement is at the level of the reference's own rounding: 1e-8 relative + 1e-11 of the response scale."""
import numpy as np
import pytest

from conftest import small_model, to_oracle

pytestmark = pytest.mark.gpu


def _check(km, x, tau2, type="UK", rtol=1e-8):
    from diceoptim_b200 import AKG
    from oracle import diceoptim as do
    ko = to_oracle(km)
    got = AKG(x, km, new_noise_var=tau2, type=type)
    ref = np.array([do.AKG(xi, ko, new_noise_var=tau2, type=type) for xi in x])
    scale = np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2)
    assert np.all(np.abs(got - ref) <= rtol * np.abs(ref) + 1e-11 * scale), (np.max(np.abs(got - ref)), ref[:5], got[:5])
    return got, ref


@pytest.mark.parametrize("trend,type", [("const", "UK"), ("const", "SK"), ("linear", "UK")])
@pytest.mark.parametrize("n", [12, 150])
def test_akg_noise_var_model(gpu_lib, trend, type, n):
    rng = np.random.default_rng(n)
    km = small_model("matern5_2", trend, n=n, d=2, seed=421, noise_var=0.02 + 0.03 * rng.random(n))
    x = rng.random((90, 2))
    got, ref = _check(km, x, tau2=0.02, type=type)
    assert np.all(got >= -1e-12) and np.max(ref) > 1e-4
    _check(km, x[:20], tau2=0.0, type=type)       # new.noise.var = 0 (default)
    km.release()


def test_akg_other_kernels_and_single_point(gpu_lib):
    from diceoptim_b200 import AKG
    rng = np.random.default_rng(3)
    for name in ("gauss", "exp", "matern3_2", "powexp"):
        km = small_model(name, "const", n=40, d=3, seed=7, noise_var=0.05)
        x = rng.random((33, 3))
        got, _ = _check(km, x, tau2=0.05)
        assert isinstance(AKG(x[0], km, new_noise_var=0.05), float)
        assert AKG(x[0], km, new_noise_var=0.05) == got[0]
        km.release()


def test_akg_nugget_and_noise_free_models_values_only(gpu_lib):
    """With a homogeneous nugget (or no noise) the reference's cn is rounding noise around mu.x; the slopes'
    order is not reproducible but the AKG value is."""
    rng = np.random.default_rng(5)
    x = rng.random((60, 2))
    kn = small_model("matern5_2", "const", n=30, d=2, seed=2, nugget=0.05)
    _check(kn, x, tau2=0.05, type="SK", rtol=1e-7)
    _check(kn, x, tau2=0.05, type="UK", rtol=1e-7)
    kd = small_model("matern5_2", "const", n=30, d=2, seed=2)
    _check(kd, x, tau2=0.01, type="SK", rtol=1e-7)
    kn.release(); kd.release()


def test_akg_zero_at_design_points_and_size_limit(gpu_lib):
    from diceoptim_b200 import AKG
    km = small_model("matern5_2", "const", n=25, d=2, seed=1)          # noise-free: sk.x = 0 at the design
    assert np.all(AKG(km.X[:5], km, new_noise_var=0.1) == 0.0)          # guard of R/AKG.R:115
    km.release()


@pytest.mark.parametrize("trend,type", [("const", "UK"), ("const", "SK"), ("linear", "UK")])
def test_akg_grad_matches_oracle(gpu_lib, trend, type):
    """AKG.grad (R/AKG.grad.R) vs the oracle's restatement (reference quirks included); the oracle itself is
    checked against finite differences of AKG for type SK in tests/test_oracle_basic.py."""
    from diceoptim_b200 import AKG_grad
    from oracle import diceoptim as do
    rng = np.random.default_rng(11)
    km = small_model("matern5_2", trend, n=90, d=3, seed=421, noise_var=0.02 + 0.03 * rng.random(90))
    ko = to_oracle(km)
    x = rng.random((40, 3))
    got = AKG_grad(x, km, new_noise_var=0.02, type=type)
    assert got.shape == (40, 3)
    for j in range(40):
        ref = do.AKG_grad(x[j], ko, new_noise_var=0.02, type=type)
        assert np.max(np.abs(got[j] - ref)) <= 1e-7 * np.max(np.abs(ref)) + 1e-13, (j, got[j], ref)
    assert np.allclose(AKG_grad(x[3], km, new_noise_var=0.02, type=type), got[3], rtol=1e-12, atol=1e-300)
    km.release()


def test_akg_grad_other_kernels_and_guard(gpu_lib):
    from diceoptim_b200 import AKG_grad
    from oracle import diceoptim as do
    rng = np.random.default_rng(2)
    for name in ("gauss", "matern3_2"):
        km = small_model(name, "const", n=40, d=2, seed=5, noise_var=0.05)
        x = rng.random((12, 2))
        got = AKG_grad(x, km, new_noise_var=0.05)
        for j in range(12):
            ref = do.AKG_grad(x[j], to_oracle(km), new_noise_var=0.05)
            assert np.max(np.abs(got[j] - ref)) <= 1e-7 * np.max(np.abs(ref)) + 1e-13
        km.release()
    kd = small_model("matern5_2", "const", n=25, d=2, seed=1)
    assert np.all(AKG_grad(kd.X[:4], kd, new_noise_var=0.1) == 0.0)     # sk.x = 0 at the design: zero gradient
    kd.release()


def test_aei_and_eqi_match_oracle(gpu_lib):
    """The noisy-criteria family of SURVEY 8(f): AEI (R/AEI.R) and EQI (R/EQI.R), default y.min / q.min included."""
    from diceoptim_b200 import AEI, EQI
    from oracle import diceoptim as do
    rng = np.random.default_rng(9)
    for type in ("UK", "SK"):
        km = small_model("matern5_2", "const", n=60, d=3, seed=421, noise_var=0.03)
        ko = to_oracle(km)
        x = np.vstack([rng.random((50, 3)), km.X[:2]])
        a = AEI(x, km, new_noise_var=0.03, type=type)
        e = EQI(x, km, new_noise_var=0.03, beta=0.9, type=type)
        ra = np.array([do.AEI(xi, ko, new_noise_var=0.03, type=type) for xi in x])
        re = np.array([do.EQI(xi, ko, new_noise_var=0.03, beta=0.9, type=type) for xi in x])
        assert np.all(np.abs(a - ra) <= 1e-10 * np.abs(ra) + 1e-12 * ra.max())
        assert np.all(np.abs(e - re) <= 1e-10 * np.abs(re) + 1e-12 * re.max())
        assert isinstance(AEI(x[0], km, new_noise_var=0.03, type=type), float)
        km.release()


@pytest.mark.parametrize("trend,type", [("const", "UK"), ("const", "SK"), ("linear", "UK")])
def test_aei_grad_and_eqi_grad_match_oracle(gpu_lib, trend, type):
    from diceoptim_b200 import AEI_grad, EQI_grad
    from oracle import diceoptim as do
    rng = np.random.default_rng(12)
    km = small_model("matern5_2", trend, n=70, d=3, seed=421, noise_var=0.03)
    ko = to_oracle(km)
    x = np.vstack([rng.random((30, 3)), km.X[:1]])
    ga = AEI_grad(x, km, new_noise_var=0.03, type=type)
    ge = EQI_grad(x, km, new_noise_var=0.03, beta=0.9, type=type)
    for j in range(x.shape[0]):
        ra = do.AEI_grad(x[j], ko, new_noise_var=0.03, type=type)
        re = do.EQI_grad(x[j], ko, new_noise_var=0.03, beta=0.9, type=type)
        assert np.max(np.abs(ga[j] - ra)) <= 1e-10 * np.max(np.abs(ra)) + 1e-14
        assert np.max(np.abs(ge[j] - re)) <= 1e-10 * np.max(np.abs(re)) + 1e-14
    km.release()


def test_akg_envir_reuse(gpu_lib):
    """AKG(x, envir) then AKG.grad(x, envir) (R/AKG.R:221-228, R/AKG.grad.R:84-94): one launch serves both calls."""
    from diceoptim_b200 import AKG, AKG_grad
    km = small_model("matern5_2", "const", n=50, d=2, seed=12, noise_var=0.05)
    x = np.random.default_rng(2).random((6, 2))
    env = {}
    v = AKG(x, km, new_noise_var=0.05, envir=env)
    g = AKG_grad(x, km, new_noise_var=0.05, envir=env)
    assert g is env["AKG.grad"]
    assert np.array_equal(g, AKG_grad(x, km, new_noise_var=0.05))
    assert np.allclose(v, AKG(x, km, new_noise_var=0.05), rtol=1e-12, atol=1e-15)
    assert AKG_grad(x[:3], km, new_noise_var=0.05, envir=env) is not env["AKG.grad"]
    km.release()
