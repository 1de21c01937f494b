"""One launch of every fused kernel on mid-size inputs, for a single `ncu --set full` pass over the family.
usage: python tools/profile_all.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
from diceoptim_b200 import _lib, synth
lib = _lib.load()
p = lambda t: C.c_void_p(t.data_ptr())
km = synth.config_model("cfg4_sphere10_n2000"); h = km.handle(0)
N = 296 * 64 * 2
x = torch.rand((km.d, N), dtype=torch.float64, device="cuda"); o = torch.empty(N, dtype=torch.float64, device="cuda")
gr = torch.empty((km.d, N), dtype=torch.float64, device="cuda")
pl = float(np.min(km.y))
for _ in range(int(os.environ.get("PROFILE_REPS", "2"))):
    _lib.check(lib.dk_ei_grad_dev(h.ptr, p(x), N, pl, 1, 1, 0, p(gr), p(o), None), "grad")
torch.cuda.synchronize(); km.release()
km3 = synth.config_model("cfg3_rosenbrock4_n500"); h3 = km3.handle(0)
B = 296 * 16 * 4
xq = torch.rand((km3.d, B * 4), dtype=torch.float64, device="cuda"); oq = torch.empty(B, dtype=torch.float64, device="cuda")
gq = torch.empty((km3.d, B * 4), dtype=torch.float64, device="cuda")
for _ in range(int(os.environ.get("PROFILE_REPS", "2"))):
    _lib.check(lib.dk_qei_dev(h3.ptr, p(xq), B, 4, 0.0, 1, 1, p(oq), None), "qei")
    _lib.check(lib.dk_qei_grad_dev(h3.ptr, p(xq), B, 4, 0.0, 1, 1, p(gq), None), "qei_grad")
torch.cuda.synchronize(); km3.release()
km5 = synth.config_model("cfg5_noisy20_n5000"); h5 = km5.handle(0)
Na = 296 * 64
xa = torch.rand((km5.d, Na), dtype=torch.float64, device="cuda"); oa = torch.empty(Na, dtype=torch.float64, device="cuda")
for _ in range(int(os.environ.get("PROFILE_REPS", "2"))):
    _lib.check(lib.dk_akg_dev(h5.ptr, p(xa), Na, 0.1, 1, p(oa), None), "akg")
torch.cuda.synchronize()
print("profile_all done")
