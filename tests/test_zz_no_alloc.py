"""GPU: the allocation-free EI entry point (dk_ei_dev_noalloc) against dk_ei_dev.

Who needs it: a host that may not allocate device memory while it runs (a serving loop with a
fixed memory budget, or code under another framework's memory pool, where a raw cudaMalloc
stalls the device and can fail).  The entry point keeps the solve's scratch in a fixed array
inside the library, which every call shares and which is never cleared, so the test calls it
twice in a row (the second call sees the first call's leftovers, with a different batch) and
checks both against the ordinary entry point, to the tolerance the other EI tests use.
This file sorts last on purpose: a run that stops at the first failure has already run the rest.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import small_model

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _both(lib, km, x_host):
    import torch
    from diceoptim_b200 import _lib
    N, d = x_host.shape
    x = torch.from_numpy(np.ascontiguousarray(x_host.T)).cuda()      # d x N row-major == N x d column-major
    outs = {}
    plugin = float(np.min(km.y))
    h = km.handle(0)
    for name in ("dk_ei_dev", "dk_ei_dev_noalloc"):
        ei, mean, sd = (torch.full((N,), float("nan"), dtype=torch.float64, device="cuda") for _ in range(3))
        _lib.check(getattr(lib, name)(h.ptr, _ptr(x), N, plugin, 1, 1, 0, _ptr(ei), _ptr(mean), _ptr(sd), None), name)
        torch.cuda.synchronize()
        outs[name] = [t.cpu().numpy() for t in (ei, mean, sd)]
    return outs


def _assert_same(outs):
    a, b = outs["dk_ei_dev"], outs["dk_ei_dev_noalloc"]
    for u, v in zip(a, b):
        assert np.array_equal(np.isnan(u), np.isnan(v))
        ok = ~np.isnan(u)
        floor = 1e-12 * np.max(np.abs(u[ok]))
        assert np.all(np.abs(u[ok] - v[ok]) <= 1e-9 * np.abs(u[ok]) + floor)


def test_noalloc_matches_regular_entry_point_on_two_calls_in_a_row(gpu_lib):
    rng = np.random.default_rng(33)
    km = small_model("matern5_2", "const", n=300, d=4, seed=31)      # 3 block rows: panels are re-read
    _assert_same(_both(gpu_lib, km, rng.random((700, 4))))
    _assert_same(_both(gpu_lib, km, rng.random((1500, 4))))          # second call, stale scratch, other batch
    km2 = small_model("gauss", "linear", n=90, d=3, seed=32)         # another handle shares the same array
    _assert_same(_both(gpu_lib, km2, rng.random((257, 3))))
    km.release(); km2.release()


def test_noalloc_allocates_nothing_and_caps_the_grid(gpu_lib):
    """The entry point's defining property: device memory in use is the same before and after a call (the ordinary entry
    point is warmed first, so neither the context nor the module load is counted), for a model whose panels only partly
    fit the fixed array (n = 2000: 1 MiB per CTA, 64 of the 296 CTAs get one), with 64- and with 16-candidate tiles."""
    import torch
    from diceoptim_b200 import synth, _lib
    km = synth.config_model("cfg4_sphere10_n2000")
    h = km.handle(0)
    rng = np.random.default_rng(34)
    for N in (6400, 300):                                  # 100 wide tiles (> 64 panels) / 19 narrow tiles
        _assert_same(_both(gpu_lib, km, rng.random((N, km.d))))         # warm both entry points
        x = torch.rand((km.d, N), dtype=torch.float64, device="cuda")
        ei = torch.empty(N, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.mem_get_info()[0]
        bytes_before = gpu_lib.dk_model_device_bytes(h.ptr)
        for _ in range(3):
            _lib.check(gpu_lib.dk_ei_dev_noalloc(h.ptr, _ptr(x), N, float(np.min(km.y)), 1, 1, 0, _ptr(ei), None, None, None), "noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == before      # free device memory unchanged: nothing allocated or freed
        assert gpu_lib.dk_model_device_bytes(h.ptr) == bytes_before
        assert torch.isfinite(ei).all()
    km.release()


def test_noalloc_argument_errors(gpu_lib):
    """Bad arguments are refused with a message and nothing runs.  (The size limit itself, a per-CTA
    panel larger than the 64 MiB array, needs n + p + 1 > 131072 rows, too large a model to build in a test.)"""
    import torch
    km = small_model("exp", "const", n=40, d=2, seed=5)
    h = km.handle(0)
    x = torch.rand((2, 10), dtype=torch.float64, device="cuda")
    out = torch.empty(10, dtype=torch.float64, device="cuda")
    assert gpu_lib.dk_ei_dev_noalloc(h.ptr, _ptr(x), 10, 0.0, 1, 1, 0, _ptr(out), None, None, None) == 0
    assert gpu_lib.dk_ei_dev_noalloc(h.ptr, _ptr(x), 10, 0.0, 1, 1, 0, None, None, None, None) != 0
    assert b"NULL" in gpu_lib.dk_last_error()
    assert gpu_lib.dk_ei_dev_noalloc(None, _ptr(x), 10, 0.0, 1, 1, 0, _ptr(out), None, None, None) != 0
    km.release()
