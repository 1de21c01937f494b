"""qEI.grad for q = 2 .. 10: kernel time per batch (moments + gradient kernels, CUDA events of the library) on the cfg3 model
(rosenbrock4, n = 500, d = 4).  usage: python tools/time_qei_grad_gen.py [B]"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from diceoptim_b200 import _lib, synth
lib = _lib.load()
km = synth.config_model("cfg3_rosenbrock4_n500"); h = km.handle(0); lib.dk_model_set_timing(h.ptr, 1)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 148
out = {}
for q in (2, 4, 5, 6, 8, 10):
    x = torch.rand((km.d, B * q), dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(q))
    g = torch.empty((km.d, B * q), dtype=torch.float64, device="cuda")
    ms = []
    for _ in range(3):
        _lib.check(lib.dk_qei_grad_dev(h.ptr, C.c_void_p(x.data_ptr()), B, q, float(np.median(km.y)), 1, 1, C.c_void_p(g.data_ptr()), None), "qei grad")
        torch.cuda.synchronize(); ms.append(lib.dk_model_last_kernel_ms(h.ptr))
    out[q] = {"batches": B, "kernel_ms": min(ms), "ms_per_batch": min(ms) / B, "finite": bool(torch.isfinite(g).all())}
print(json.dumps(out))
