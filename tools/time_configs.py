"""Wall-clock of the host-facing calls (host buffers, copies included) on BASELINE configs 1-3 and 5.
usage: python tools/time_configs.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from diceoptim_b200 import AKG, EI, EI_grad, kriging_quantile, qEI, qEI_grad, synth


def t(f, reps=3):
    f(); best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); f(); best = min(best, time.perf_counter() - t0)
    return best


km = synth.config_model("cfg1_branin2_n20"); x = synth.config_candidates("cfg1_branin2_n20")
print("cfg1 n=20 N=1e4: EI %.2f ms, EI.grad %.2f ms" % (1e3 * t(lambda: EI(x, km)), 1e3 * t(lambda: EI_grad(x, km))))
km = synth.config_model("cfg2_hartman4_n200"); x = synth.config_candidates("cfg2_hartman4_n200")
dt = t(lambda: EI(x, km)); print("cfg2 n=200 N=1e6: EI %.2f ms = %.2e cand/s" % (1e3 * dt, x.shape[0] / dt))
km = synth.config_model("cfg3_rosenbrock4_n500"); xb = np.random.default_rng(13).random((100000, 4, 4))
dt = t(lambda: qEI(xb, km, plugin=float(np.median(km.y)))); print("cfg3 n=500 B=1e5 q=4: qEI %.2f ms = %.2e evals/s" % (1e3 * dt, 1e5 / dt))
dt = t(lambda: qEI_grad(xb, km, plugin=float(np.median(km.y)))); print("cfg3 qEI.grad %.2f ms = %.2e evals/s" % (1e3 * dt, 1e5 / dt))
km = synth.config_model("cfg5_noisy20_n5000"); x = synth.config_candidates("cfg5_noisy20_n5000", N=296 * 64 * 4)
dt = t(lambda: AKG(x, km, new_noise_var=0.1), 2); print("cfg5 n=5000 N=%d: AKG %.1f ms = %.2e cand/s" % (x.shape[0], 1e3 * dt, x.shape[0] / dt))
dt = t(lambda: kriging_quantile(x, km, beta=0.1), 2); print("cfg5 quantile %.1f ms = %.2e cand/s" % (1e3 * dt, x.shape[0] / dt))
