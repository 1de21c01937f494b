"""Device-resident EI on the small-n BASELINE configs (cfg1 n = 20 on the 1e4 grid, cfg2 n = 200 on 1e6 candidates):
kernel time by the library's CUDA events, algorithmic HBM bytes (8d read + 8 written per candidate) and flops.
usage: python tools/profile_small.py"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from diceoptim_b200 import _lib, synth
lib = _lib.load()
out = {}
for key in ("cfg1_branin2_n20", "cfg2_hartman4_n200"):
    km = synth.config_model(key); h = km.handle(0); lib.dk_model_set_timing(h.ptr, 1)
    xh = synth.config_candidates(key); N = xh.shape[0]
    x = torch.from_numpy(np.ascontiguousarray(xh.T)).cuda()  # d x N row-major = N x d column-major
    ei = torch.empty(N, dtype=torch.float64, device="cuda"); ms = []
    for _ in range(5):
        _lib.check(lib.dk_ei_dev(h.ptr, C.c_void_p(x.data_ptr()), N, float(np.min(km.y)), 1, 1, 0, C.c_void_p(ei.data_ptr()), None, None, None), "ei")
        torch.cuda.synchronize(); ms.append(lib.dk_model_last_kernel_ms(h.ptr))
    t = min(ms[1:]) * 1e-3; n, d = km.n, km.d
    out[key] = {"n": n, "d": d, "N": N, "kernel_ms": t * 1e3, "candidates_per_s": N / t,
                "algorithmic_GBps": N * (8 * d + 8) / t * 1e-9, "algorithmic_TFLOPs": N * (n * n + 4 * n + 2 * n) / t * 1e-12,
                "covgen_factors_per_s": N * n * d / t}
    km.release()
print(json.dumps(out))
