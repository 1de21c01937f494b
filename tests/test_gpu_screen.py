"""dk_screen: any criterion, candidates given or generated on the device (include/dk_b200.h).  The generator is
checked against its numpy restatement (tests/philox_ref.py) and the screen over generated candidates against the same
screen over that matrix passed from the host, bit for bit."""
import ctypes as C

import numpy as np
import pytest

from conftest import small_model, to_oracle
from philox_ref import uniform_rows

pytestmark = pytest.mark.gpu


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _specs(d, seed, lower, upper, criterion, **kw):
    from diceoptim_b200 import _lib
    lo = np.asarray(lower, float); up = np.asarray(upper, float)
    gen = _lib.dk_uniform_spec(seed=seed, lower=lo.ctypes.data_as(_lib.c_double_p), upper=up.ctypes.data_as(_lib.c_double_p))
    crit = _lib.dk_crit_spec(criterion=criterion, type=1, minimization=1, proxy=0, plugin=kw.get("plugin", 0.0),
                             qnorm_beta=kw.get("qnorm_beta", 0.0), new_noise_var=kw.get("tau2", 0.0))
    return gen, crit, (lo, up)


@pytest.mark.parametrize("criterion", [0, 1, 2])
def test_screen_generated_equals_screen_given(gpu_lib, criterion):
    from diceoptim_b200 import _lib
    km = small_model("matern5_2", "const", n=90, d=3, seed=31, noise_var=0.03)
    h = km.handle(0)
    N, k, seed = 50_001, 6, 0x1234_5678_9ABC_DEF1
    lower, upper = [0.1, 0.0, -0.5], [0.9, 1.0, 0.5]
    gen, crit, keep = _specs(3, seed, lower, upper, criterion, plugin=float(np.min(km.y)), qnorm_beta=-1.2815515655446004, tau2=0.03)
    ia = np.empty(k, np.int64); va = np.empty(k); xa = np.empty((k, 3), order="F")
    _lib.check(gpu_lib.dk_screen(h.ptr, None, 0, C.byref(gen), N, C.byref(crit), k, _vp(ia), _vp(va), _vp(xa), None), "dk_screen(gen)")
    x = np.asfortranarray(uniform_rows(seed, 0, N, 3, lower, upper))
    assert x.min() >= -0.5 and x.max() <= 1.0
    assert np.array_equal(xa, x[ia])                       # the generator, bit for bit
    ib = np.empty(k, np.int64); vb = np.empty(k)
    _lib.check(gpu_lib.dk_screen(h.ptr, _vp(x), 0, None, N, C.byref(crit), k, _vp(ib), _vp(vb), None, None), "dk_screen(x)")
    assert np.array_equal(ia, ib) and np.array_equal(va, vb)
    # against the oracle's ranking of the same matrix
    from oracle import diceoptim as do
    ko = to_oracle(km)
    if criterion == 0:
        ref = do.EI(x, ko)
        assert ia[0] == do.which_max(ref)
    elif criterion == 1:
        ref = do.kriging_quantile(x, ko, beta=0.1)
        assert ia[0] == int(np.argmin(ref)) and np.all(np.diff(va) >= 0)
    else:
        sub = np.concatenate([ia, np.arange(0, N, 997)])
        ref = np.array([do.AKG(x[i], ko, new_noise_var=0.03) for i in sub])
        assert np.argmax(ref) == 0 or abs(ref.max() - ref[0]) <= 1e-8 * abs(ref[0])
    km.release()


def test_multi_screen_generated_is_independent_of_device_count(gpu_lib):
    from diceoptim_b200 import _lib
    from diceoptim_b200.km import MultiDeviceModel
    km = small_model("gauss", "const", n=70, d=2, seed=32)
    N, k, seed = 120_000, 4, 77
    gen, crit, keep = _specs(2, seed, [0, 0], [1, 1], 0, plugin=float(np.min(km.y)))
    res = []
    for devices in ([0], [0, 0, 0]):
        mdm = MultiDeviceModel(km, devices)
        idx = np.empty(k, np.int64); val = np.empty(k); xs = np.empty((k, 2), order="F")
        _lib.check(gpu_lib.dk_multi_screen(mdm.ptr, None, C.byref(gen), N, C.byref(crit), k, _vp(idx), _vp(val), _vp(xs)), "dk_multi_screen")
        res.append((idx.copy(), val.copy(), xs.copy()))
        mdm.close()
    for a, b in zip(res[0], res[1]):
        assert np.array_equal(a, b)
    assert np.array_equal(res[0][2], uniform_rows(seed, 0, N, 2)[res[0][0]])
    km.release()
