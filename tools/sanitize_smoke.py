"""Smallest workload that touches every kernel of the library once (for compute-sanitizer, one tool per run):
   compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from diceoptim_b200 import (AEI, AKG, AKG_grad, EI, EI_grad, EQI, kriging_quantile, kriging_quantile_grad, qEI, qEI_grad,
                            sampleFromEI, screen_EI, synth)
rng = np.random.default_rng(0)
X = synth.lhs(150, 3, 1); y = np.sin(4 * X.sum(1))
for kw in (dict(trend="const", noise_var=0.02), dict(trend="linear")):
    km = synth.make_km(X, y, "matern5_2", 0.5, float(np.var(y)), **kw)
    x = rng.random((200, 3))
    out = [EI(x, km), EI_grad(x, km), kriging_quantile(x, km), kriging_quantile_grad(x[:70], km), screen_EI(x, km, k=3)[1],
           AKG(x, km, new_noise_var=0.02), AKG_grad(x[:70], km, new_noise_var=0.02), AEI(x, km, new_noise_var=0.02),
           EQI(x, km, new_noise_var=0.02), qEI(rng.random((40, 4, 3)), km, plugin=0.0), qEI_grad(rng.random((20, 3, 3)), km, plugin=0.0),
           sampleFromEI(km, n=3, seed=1)]
    print(kw, [float(np.nansum(o)) for o in out])
    km.release()
print("sanitize smoke done")
