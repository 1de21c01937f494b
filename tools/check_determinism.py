"""Race detector for the fused kernels: evaluates the same candidates as one call and as shards, several times,
and demands bit-identical EI (small n makes the window between a panel store and its bulk-copy read shortest).
usage: python tools/check_determinism.py [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from diceoptim_b200 import EI, EI_grad, synth
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
tot = 0
for cfg, N in (("cfg2_hartman4_n200", 1_000_000), ("cfg3_rosenbrock4_n500", 400_000), ("cfg4_sphere10_n2000", 200_000)):
    km = synth.config_model(cfg); x = synth.config_candidates(cfg, N=N)
    ei = EI(x, km)
    q = N // 4
    for rep in range(reps):
        lo = q * (rep % 4); xs = np.asfortranarray(x[lo:lo + q])
        bad = int(np.count_nonzero(EI(xs, km) != ei[lo:lo + q])); tot += bad
        if bad: print(cfg, "rep", rep, "mismatches", bad)
    g1 = EI_grad(x[:20000], km); g2 = EI_grad(x[:20000], km)
    gb = int(np.count_nonzero(g1 != g2)); tot += gb
    print(cfg, "done; gradient repeat mismatches", gb)
    km.release()
print("TOTAL MISMATCHES", tot)
sys.exit(1 if tot else 0)
