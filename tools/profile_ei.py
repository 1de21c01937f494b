"""Short driver for ncu: a few launches of the fused kernels on the cfg4 model (n=2000, d=10).
usage: python tools/profile_ei.py [N] [reps] [ei|grad|akg]   (DK_CFG selects the synth config)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as g

g.build()
from diceoptim_b200 import _lib, synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
mode = sys.argv[3] if len(sys.argv) > 3 else "ei"
grad = mode == "grad"
cfg = os.environ.get("DK_CFG", "cfg4_sphere10_n2000")
lib = _lib.load()
km = synth.config_model(cfg)
h = km.handle(0)
lib.dk_model_set_timing(h.ptr, 1)
x = torch.rand((km.d, N), dtype=torch.float64, device="cuda")
ei = torch.empty(N, dtype=torch.float64, device="cuda")
gr = torch.empty((km.d, N), dtype=torch.float64, device="cuda") if grad else None
plugin = float(np.min(km.y))
ms = []
for _ in range(reps):
    if mode == "akg":
        _lib.check(lib.dk_akg_dev(h.ptr, C.c_void_p(x.data_ptr()), N, 0.1, 1, C.c_void_p(ei.data_ptr()), None), "akg")
    elif grad:
        _lib.check(lib.dk_ei_grad_dev(h.ptr, C.c_void_p(x.data_ptr()), N, plugin, 1, 1, 0, C.c_void_p(gr.data_ptr()),
                                      C.c_void_p(ei.data_ptr()), None), "grad")
    else:
        _lib.check(lib.dk_ei_dev(h.ptr, C.c_void_p(x.data_ptr()), N, plugin, 1, 1, 0, C.c_void_p(ei.data_ptr()), None, None,
                                 None), "ei")
    torch.cuda.synchronize()
    ms.append(lib.dk_model_last_kernel_ms(h.ptr))
fl = (km.n ** 2 + 4 * km.n + 2 * km.n * km.p) * (2 if grad else 1)
print({"cfg": cfg, "N": N, "grad": grad, "kernel_ms": ms, "cand_per_s": N / (min(ms) * 1e-3),
       "tflops_alg": fl * N / (min(ms) * 1e-3) * 1e-12})
