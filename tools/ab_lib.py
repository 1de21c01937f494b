"""Time the cfg4 EI (or EI.grad) kernel with an alternative build of the library: python tools/ab_lib.py <lib.so> [N] [ei|grad|qei|qeigrad]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from diceoptim_b200 import _lib, synth
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
lib = _lib.load()
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
mode = sys.argv[3] if len(sys.argv) > 3 else "ei"
if mode in ("qei", "qeigrad"):
    km = synth.config_model("cfg3_rosenbrock4_n500"); h = km.handle(0); q = 4
    x = torch.rand((km.d, N * q), dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(3))
    o = torch.empty(N * (q * km.d if mode == "qeigrad" else 1), dtype=torch.float64, device="cuda")
    pl = float(np.median(km.y)); ms = []
    for _ in range(4):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
        if mode == "qei": _lib.check(lib.dk_qei_dev(h.ptr, C.c_void_p(x.data_ptr()), N, q, pl, 1, 1, C.c_void_p(o.data_ptr()), None), "qei")
        else: _lib.check(lib.dk_qei_grad_dev(h.ptr, C.c_void_p(x.data_ptr()), N, q, pl, 1, 1, C.c_void_p(o.data_ptr()), None), "qeigrad")
        e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print(os.path.basename(sys.argv[1]), mode, N, [round(m, 2) for m in ms], "evals/s %.3e" % (N / (min(ms) * 1e-3)), "checksum %.15e" % float(o.sum()))
    sys.exit(0)
km = synth.config_model(os.environ.get("DK_CFG", "cfg4_sphere10_n2000"))
h = km.handle(0); lib.dk_model_set_timing(h.ptr, 1)
x = torch.rand((km.d, N), dtype=torch.float64, device="cuda"); ei = torch.empty(N, dtype=torch.float64, device="cuda")
gr = torch.empty((km.d, N), dtype=torch.float64, device="cuda") if mode == "grad" else None
ms = []
for _ in range(3):
    if mode == "grad":
        _lib.check(lib.dk_ei_grad_dev(h.ptr, C.c_void_p(x.data_ptr()), N, float(np.min(km.y)), 1, 1, 0, C.c_void_p(gr.data_ptr()), C.c_void_p(ei.data_ptr()), None), "grad")
    else:
        _lib.check(lib.dk_ei_dev(h.ptr, C.c_void_p(x.data_ptr()), N, float(np.min(km.y)), 1, 1, 0, C.c_void_p(ei.data_ptr()), None, None, None), "ei")
    torch.cuda.synchronize(); ms.append(lib.dk_model_last_kernel_ms(h.ptr))
print(os.path.basename(sys.argv[1]), N, [round(m, 2) for m in ms], "cand/s %.3e" % (N / (min(ms) * 1e-3)), "checksum", float(ei.sum()))
