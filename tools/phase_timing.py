"""Per-phase cycle accounting of k_predict_fused_tma (thread 0 of every CTA, clock64 between phase marks).
Builds a separate library with -DDK_PHASE_TIMING, runs the cfg4 workload once and prints the cycle shares.
usage: python tools/phase_timing.py [N]"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import __graft_entry__ as g

lib_t = os.path.join(ROOT, "diceoptim_b200", "libdk_b200_timing.so")
srcs = [os.path.join(g.CSRC, f) for f in os.listdir(g.CSRC) if f.endswith(".cu")]
subprocess.run([os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")] + g.NVCC_FLAGS + ["-DDK_PHASE_TIMING", "-o", lib_t] + srcs + ["-lgomp"], check=True)
from diceoptim_b200 import _lib, synth
_lib.LIB_PATH = lib_t
lib = _lib.load()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 592000
km = synth.config_model(os.environ.get("DK_CFG", "cfg4_sphere10_n2000"))
h = km.handle(0)
x = torch.rand((km.d, N), dtype=torch.float64, device="cuda")
ei = torch.empty(N, dtype=torch.float64, device="cuda")
for _ in range(2):
    _lib.check(lib.dk_ei_dev(h.ptr, C.c_void_p(x.data_ptr()), N, float(np.min(km.y)), 1, 1, 0, C.c_void_p(ei.data_ptr()), None, None, None), "ei")
torch.cuda.synchronize()
# read the device symbol through the CUDA runtime of the library's own context
rt = C.CDLL("libcudart.so")
names = ["P1+P2 covgen", "-", "P3 GEMM", "P4 diag", "P5 panel store", "s2 + barrier", "loop top", "return from solve"]
buf = (C.c_ulonglong * 8)()
lib.dk_debug_phase_cycles.argtypes = [C.c_void_p]
lib.dk_debug_phase_cycles(buf)
tot = sum(buf)
for nm, v in zip(names, buf):
    print(f"{nm:20s} {v:16d} {100.0 * v / max(tot, 1):6.2f} %")
