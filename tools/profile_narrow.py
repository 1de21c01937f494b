"""Short driver for ncu: the narrow-tile kernels on an optimiser-sized batch (cfg4 model, N candidates, default 16 x 19 = 304).
usage: python tools/profile_narrow.py [N] [reps]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
from diceoptim_b200 import _lib, synth
N = int(sys.argv[1]) if len(sys.argv) > 1 else 304
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lib = _lib.load()
km = synth.config_model(os.environ.get("DK_CFG", "cfg4_sphere10_n2000")); h = km.handle(0)
lib.dk_model_set_timing(h.ptr, 1)
lib.dk_model_set_tile_mode(h.ptr, 2)
x = torch.rand((km.d, N), dtype=torch.float64, device="cuda"); ei = torch.empty(N, dtype=torch.float64, device="cuda")
gr = torch.empty((km.d, N), dtype=torch.float64, device="cuda")
p = lambda t: C.c_void_p(t.data_ptr())
pl = float(np.min(km.y)); out = {}
for name in ("ei", "grad"):
    ms = []
    for _ in range(reps):
        if name == "ei": _lib.check(lib.dk_ei_dev(h.ptr, p(x), N, pl, 1, 1, 0, p(ei), None, None, None), name)
        else: _lib.check(lib.dk_ei_grad_dev(h.ptr, p(x), N, pl, 1, 1, 0, p(gr), p(ei), None), name)
        torch.cuda.synchronize(); ms.append(lib.dk_model_last_kernel_ms(h.ptr))
    out[name + "_kernel_ms"] = ms
print({"N": N, **out})
