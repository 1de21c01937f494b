"""Wall-clock of the multistart optimisers: batched polish (one launch per iteration for all restarts) vs the
one-restart-at-a-time scipy path that mirrors the reference's optim() loop.  usage: python tools/time_multistart.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from diceoptim_b200 import max_crit, max_EI, max_qEI, synth

km = synth.config_model("cfg4_sphere10_n2000"); d = km.d
lo, up = np.zeros(d), np.ones(d)
max_crit(km, lower=lo, upper=up, control={"restarts": 2, "n.ei": 100}, seed=0)    # warm-up
for opt in ("batched", "scipy"):
    t0 = time.perf_counter()
    r = max_crit(km, lower=lo, upper=up, control={"restarts": 10, "n.ei": 5000, "optimizer": opt}, seed=1)
    print("cfg4 max_crit restarts=10 n.ei=5000 %-7s: %.3f s  value %.6e" % (opt, time.perf_counter() - t0, r["value"]))
t0 = time.perf_counter()
r = max_EI(km, lower=lo, upper=up, control={"n.screen": 1 << 20, "pop.size": 320}, seed=1)
print("cfg4 max_EI screen 2^20 + 320 polished starts: %.3f s  value %.6e" % (time.perf_counter() - t0, r["value"]))
km.release()
km = synth.config_model("cfg3_rosenbrock4_n500"); d = km.d
lo, up = np.zeros(d), np.ones(d)
par0 = np.random.default_rng(0).random((16, 4, d))
max_qEI(km, 4, lo, up, optimcontrol={"parinit": par0[:2], "maxit": 2})
for opt in ("batched", "scipy"):
    t0 = time.perf_counter()
    r = max_qEI(km, 4, lo, up, optimcontrol={"parinit": par0, "maxit": 50, "optimizer": opt})
    print("cfg3 max_qEI q=4 nStarts=16 maxit=50 %-7s: %.3f s  value %.6e" % (opt, time.perf_counter() - t0, r["value"]))
