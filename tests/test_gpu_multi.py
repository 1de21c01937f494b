"""dk_multi_*: one model on several GPUs of one process (include/dk_b200.h).  The tests run with however many devices
are visible (1 on the default GPU box, 2/4/8 under `gpurun --gpus N`) and, independently of that, with several
replicas on device 0 (dk_model_clone), which exercises the same split / merge code.  Results must equal the
single-device entry points BIT FOR BIT: every candidate goes through the same kernel, only the range it sits in
changes.  Also covers the chunked host pipeline on ranges that are not multiples of the chunk or tile size."""
import ctypes as C

import numpy as np
import pytest

from conftest import small_model, to_oracle

pytestmark = pytest.mark.gpu


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _desc(km, device=0):
    from diceoptim_b200.km import make_desc
    return make_desc(km, device)


def _multi(gpu_lib, km, ndev=0, devices=None):
    desc, keep = _desc(km)
    mm = C.c_void_p()
    arr = None if devices is None else (C.c_int * len(devices))(*devices)
    rc = gpu_lib.dk_multi_create(C.byref(desc), arr, ndev if devices is None else len(devices), C.byref(mm))
    assert rc == 0, gpu_lib.dk_last_error()
    return mm, keep


def _check_all(gpu_lib, km, mm, N, seed=0):
    """every dk_multi_* entry against its single-device counterpart, bit for bit"""
    from diceoptim_b200 import _lib
    d = km.d
    x = np.asfortranarray(np.random.default_rng(seed).random((N, d)))
    h = km.handle(0)
    pl = float(np.min(km.y))
    a = [np.empty(N) for _ in range(3)]; b = [np.empty(N) for _ in range(3)]
    _lib.check(gpu_lib.dk_ei(h.ptr, _vp(x), N, pl, 1, 1, 0, *map(_vp, a)), "dk_ei")
    _lib.check(gpu_lib.dk_multi_ei(mm, _vp(x), N, pl, 1, 1, 0, *map(_vp, b)), "dk_multi_ei")
    for u, v in zip(a, b):
        assert np.array_equal(u, v, equal_nan=True)
    ga = np.empty((N, d), order="F"); gb = np.empty((N, d), order="F"); ea = np.empty(N); eb = np.empty(N)
    _lib.check(gpu_lib.dk_ei_grad(h.ptr, _vp(x), N, pl, 1, 1, 0, _vp(ga), _vp(ea)), "dk_ei_grad")
    _lib.check(gpu_lib.dk_multi_ei_grad(mm, _vp(x), N, pl, 1, 1, 0, _vp(gb), _vp(eb)), "dk_multi_ei_grad")
    assert np.array_equal(ga, gb, equal_nan=True) and np.array_equal(ea, eb, equal_nan=True)
    _lib.check(gpu_lib.dk_quantile(h.ptr, _vp(x), N, -1.28, 1, _vp(ea), _vp(ga)), "dk_quantile")
    _lib.check(gpu_lib.dk_multi_quantile(mm, _vp(x), N, -1.28, 1, _vp(eb), _vp(gb)), "dk_multi_quantile")
    assert np.array_equal(ga, gb) and np.array_equal(ea, eb)
    _lib.check(gpu_lib.dk_predict(h.ptr, _vp(x), N, 0, _vp(a[0]), _vp(a[1])), "dk_predict")
    _lib.check(gpu_lib.dk_multi_predict(mm, _vp(x), N, 0, _vp(b[0]), _vp(b[1])), "dk_multi_predict")
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    if km.noise_var is not None:
        _lib.check(gpu_lib.dk_akg(h.ptr, _vp(x), N, 0.05, 1, _vp(ea)), "dk_akg")
        _lib.check(gpu_lib.dk_multi_akg(mm, _vp(x), N, 0.05, 1, _vp(eb)), "dk_multi_akg")
        assert np.array_equal(ea, eb, equal_nan=True)
    k = 5
    ia = np.empty(k, np.int64); ib = np.empty(k, np.int64); va = np.empty(k); vb = np.empty(k)
    _lib.check(gpu_lib.dk_screen_ei(h.ptr, _vp(x), 0, N, pl, 1, 1, 0, k, _vp(ia), _vp(va), None), "dk_screen_ei")
    _lib.check(gpu_lib.dk_multi_screen_ei(mm, _vp(x), N, pl, 1, 1, 0, k, _vp(ib), _vp(vb)), "dk_multi_screen_ei")
    assert np.array_equal(ia, ib) and np.array_equal(va, vb)
    q = 3
    B = N // q
    xb = np.asfortranarray(x[:B * q])
    qa = np.empty(B); qb = np.empty(B); gqa = np.empty((B * q, d), order="F"); gqb = np.empty((B * q, d), order="F")
    plq = float(np.median(km.y))
    _lib.check(gpu_lib.dk_qei(h.ptr, _vp(xb), B, q, plq, 1, 1, _vp(qa)), "dk_qei")
    _lib.check(gpu_lib.dk_qei_grad(h.ptr, _vp(xb), B, q, plq, 1, 1, _vp(gqa)), "dk_qei_grad")
    _lib.check(gpu_lib.dk_multi_qei(mm, _vp(xb), B, q, plq, 1, 1, _vp(qb), _vp(gqb)), "dk_multi_qei")
    assert np.array_equal(qa, qb, equal_nan=True) and np.array_equal(gqa, gqb, equal_nan=True)


def test_multi_all_visible_devices(gpu_lib):
    km = small_model("matern5_2", "const", n=150, d=3, seed=21, noise_var=0.02)
    mm, keep = _multi(gpu_lib, km, 0)
    import torch
    assert gpu_lib.dk_multi_count(mm) == torch.cuda.device_count()
    for N in (1, 63, 1000, 4099):
        _check_all(gpu_lib, km, mm, N, seed=N)
    gpu_lib.dk_multi_destroy(mm)
    km.release()


def test_multi_split_and_merge_with_replicas_on_one_device(gpu_lib):
    """three replicas of the model on device 0 (an ordinal may be listed more than once): the range split, the strided
    gathers / scatters and the top-k merge run exactly as on three GPUs, whatever the box has"""
    km = small_model("gauss", "linear", n=200, d=4, seed=22, noise_var=0.05)
    desc, keep = _desc(km)
    devs = (C.c_int * 3)(0, 0, 0)
    mm = C.c_void_p()
    assert gpu_lib.dk_multi_create(C.byref(desc), devs, 3, C.byref(mm)) == 0, gpu_lib.dk_last_error()
    assert gpu_lib.dk_multi_count(mm) == 3
    for N in (2, 64, 3001):
        _check_all(gpu_lib, km, mm, N, seed=N)
    gpu_lib.dk_multi_destroy(mm)
    km.release()


def test_multi_screen_ties_and_nan(gpu_lib):
    """duplicated candidates straddling the split: the lower global index must win, as in which.max"""
    km = small_model("matern5_2", "const", n=80, d=2, seed=23)
    mm, keep = _multi(gpu_lib, km, devices=[0, 0])
    from diceoptim_b200 import _lib
    from oracle import diceoptim as do
    N = 2000
    x = np.random.default_rng(9).random((N // 2, 2))
    x = np.asfortranarray(np.vstack([x, x]))            # every candidate appears twice, N/2 apart
    ib = np.empty(4, np.int64); vb = np.empty(4)
    _lib.check(gpu_lib.dk_multi_screen_ei(mm, _vp(x), N, float(np.min(km.y)), 1, 1, 0, 4, _vp(ib), _vp(vb)), "screen")
    ei = do.EI(x, to_oracle(km))
    assert ib[0] == do.which_max(ei) and ib[0] < N // 2
    assert ib[1] == ib[0] + N // 2 and vb[0] == vb[1]
    gpu_lib.dk_multi_destroy(mm)
    km.release()


def test_chunked_pipeline_many_chunks(gpu_lib):
    """a call several chunks long (small d => the 24 MB chunk is ~1.5e6 candidates at d = 2): the chunked host path must
    equal one device-resident launch bit for bit, with pageable memory and a ragged last chunk"""
    import torch
    km = small_model("matern5_2", "const", n=40, d=2, seed=24)
    from diceoptim_b200 import _lib
    h = km.handle(0)
    N = 3_400_001
    x = np.asfortranarray(np.random.default_rng(1).random((N, 2)))
    pl = float(np.min(km.y))
    a = np.empty(N)
    _lib.check(gpu_lib.dk_ei(h.ptr, _vp(x), N, pl, 1, 1, 0, _vp(a), None, None), "dk_ei")
    xd = torch.from_numpy(np.ascontiguousarray(x.T)).cuda()        # d x N row-major == N x d column-major
    od = torch.empty(N, dtype=torch.float64, device="cuda")
    _lib.check(gpu_lib.dk_ei_dev(h.ptr, C.c_void_p(xd.data_ptr()), N, pl, 1, 1, 0, C.c_void_p(od.data_ptr()), None, None, None), "dk_ei_dev")
    torch.cuda.synchronize()
    assert np.array_equal(a, od.cpu().numpy())
    idx = np.empty(1, np.int64); val = np.empty(1)
    _lib.check(gpu_lib.dk_screen_ei(h.ptr, _vp(x), 0, N, pl, 1, 1, 0, 1, _vp(idx), _vp(val), None), "dk_screen_ei")
    assert idx[0] == int(np.argmax(a)) and val[0] == a[idx[0]]
    km.release()


def test_multi_append_grows_every_replica(gpu_lib):
    """dk_multi_append: every replica grows in place; the multi-device EI of the grown model equals a fresh handle's"""
    from diceoptim_b200 import EI, _lib
    from diceoptim_b200.km import MultiDeviceModel
    from diceoptim_b200.update import update_km
    km = small_model("matern5_2", "const", n=120, d=3, seed=26, noise_var=0.01)
    mdm = MultiDeviceModel(km, [0, 0])
    pts = np.random.default_rng(8).random((10, 3))                       # 120 -> 130: crosses the 128-row block boundary
    grown = km
    for i, p in enumerate(pts):
        pc = np.ascontiguousarray(p)
        _lib.check(gpu_lib.dk_multi_append(mdm.ptr, _vp(pc), float(i) / 10, 0.0), "dk_multi_append")
        grown = update_km(grown, p, float(i) / 10, reuse_handles=False)
    assert gpu_lib.dk_model_n(gpu_lib.dk_multi_model(mdm.ptr, 1)) == 130
    x = np.asfortranarray(np.random.default_rng(9).random((500, 3)))
    a = np.empty(500)
    pl = float(np.median(km.y))
    _lib.check(gpu_lib.dk_multi_ei(mdm.ptr, _vp(x), 500, pl, 1, 1, 0, _vp(a), None, None), "dk_multi_ei")
    b = EI(x, grown, plugin=pl)
    assert np.all(np.abs(a - b) <= 1e-9 * np.abs(b) + 1e-12 * b.max())
    mdm.close(); grown.release(); km.release()
