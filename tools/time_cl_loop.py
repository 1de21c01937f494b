"""Constant-liar loop of max_qEI (R/max_qEI.R:288-304) at n = 2000, q = 8: model refresh per added point by
dk_model_append (in place) against a from-scratch dk_model_create of the grown model.
usage: python tools/time_cl_loop.py  -> one JSON line"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as g
g.build()
from diceoptim_b200 import synth, EI, EI_grad
from diceoptim_b200.update import update_km

km = synth.config_model("cfg4_sphere10_n2000")
x = synth.config_candidates("cfg4_sphere10_n2000", N=4096)
pts = np.random.default_rng(1).random((8, km.d))
EI(x, km); EI_grad(x[:16], km)
res = {}
# the C call alone (no Python KM bookkeeping): dk_model_append on a replica, 8 points
import ctypes as C
from diceoptim_b200 import _lib
lib = _lib.load()
hc = km.handle(0).clone()
t0 = time.perf_counter()
for i in range(8):
    xc = np.ascontiguousarray(pts[i])
    _lib.check(lib.dk_model_append(hc.ptr, C.c_void_p(xc.ctypes.data), float(np.min(km.y)), 0.0), "append")
res["dk_model_append_ms_per_point"] = 1e3 * (time.perf_counter() - t0) / 8
hc.close()
t0 = time.perf_counter()
for i in range(3):
    hh = type(km.handle(0))(km, 0); hh.close()
res["dk_model_create_ms"] = 1e3 * (time.perf_counter() - t0) / 3
for mode in ("in_place", "rebuild"):
    cur = km
    t_upd = t_fwd = t_grad = 0.0
    for i in range(8):
        t0 = time.perf_counter()
        nxt = update_km(cur, pts[i], float(np.min(km.y)), reuse_handles=("clone" if cur is km else True) if mode == "in_place" else False)
        nxt.handle(0)                                   # rebuild mode: the new handle is created here
        t1 = time.perf_counter()
        EI(x, nxt)                                      # forward path (screen)
        t2 = time.perf_counter()
        EI_grad(x[:16], nxt)                            # gradient path (backward tiles rebuilt lazily)
        t3 = time.perf_counter()
        t_upd += t1 - t0; t_fwd += t2 - t1; t_grad += t3 - t2
        if cur is not km:
            cur.release()
        cur = nxt
    chk = float(EI(x[:64], cur).sum())
    cur.release()
    res[mode] = {"refresh_ms_per_point": 1e3 * t_upd / 8, "first_screen_ms": 1e3 * t_fwd / 8, "first_grad_ms": 1e3 * t_grad / 8, "checksum": chk}
print(json.dumps({"n": km.n, "q": 8, **res}))
