import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
from diceoptim_b200 import synth, EI
for key in ("cfg3_rosenbrock4_n500", "cfg4_sphere10_n2000", "cfg5_noisy20_n5000"):
    t0 = time.perf_counter(); km = synth.config_model(key); t1 = time.perf_counter()
    h = km.handle(0); t2 = time.perf_counter()
    km.release(); h2 = km.handle(0); t3 = time.perf_counter()
    x = np.random.default_rng(0).random((64, km.d)); EI(x, km); t4 = time.perf_counter()
    print(key, "host chol %.2fs  first create %.3fs  second create %.3fs  first EI call %.3fs" % (t1 - t0, t2 - t1, t3 - t2, t4 - t3))
    km.release()
