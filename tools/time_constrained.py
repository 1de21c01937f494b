"""Wall-clock of the constrained criteria through the host-facing calls (host buffers, copies included) next to the
oracle's one-point-at-a-time restatement on a few points.  usage: python tools/time_constrained.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from diceoptim_b200 import checkPredict, crit_EFI, crit_SUR_cst, synth
from conftest import to_oracle
from oracle import diceoptim as do

n, d = 500, 4
X = synth.lhs(n, d, 3)
yo = synth.rosenbrock4(X); yo = (yo - yo.mean()) / yo.std()
yc1 = X[:, 0] + X[:, 1] - 1.1 + 0.2 * np.cos(5 * X[:, 2]); yc2 = 0.5 - X[:, 2] * X[:, 3]
nv = np.full(n, 1e-3)
fun = synth.make_km(X, yo, "matern5_2", np.full(d, 0.5), 1.0, noise_var=nv)
c1 = synth.make_km(X, yc1, "matern5_2", np.full(d, 0.6), float(np.var(yc1)), noise_var=nv)
c2 = synth.make_km(X, yc2, "matern3_2", np.full(d, 0.6), float(np.var(yc2)), noise_var=nv)
rng = np.random.default_rng(0)


def best(f, reps=3):
    f(); t = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); f(); t = min(t, time.perf_counter() - t0)
    return t


x = rng.random((100000, d))
t = best(lambda: crit_EFI(x, fun, [c1, c2]))
print("crit_EFI n=500 d=4, 2 constraints, 1e5 candidates (covdist gate included): %.1f ms = %.3g cand/s" % (1e3 * t, x.shape[0] / t))
t = best(lambda: checkPredict(x, [c1, c2, fun]))
print("checkPredict covdist, 3 models, 1e5 candidates: %.1f ms" % (1e3 * t))
fo, co = to_oracle(fun), [to_oracle(c1), to_oracle(c2)]
t0 = time.perf_counter(); [do.crit_EFI(xi, fo, co) for xi in x[:20]]; tc = (time.perf_counter() - t0) / 20
print("oracle crit_EFI: %.2f ms per point -> %.3g cand/s" % (1e3 * tc, 1 / tc))
P = rng.random((2000, d)); xs = rng.random((4096, d))
cc = {"integration.points": P}
t = best(lambda: crit_SUR_cst(xs, fun, [c1, c2], critcontrol=cc))
print("crit_SUR_cst n=500, 2 constraints, M=2000 integration points, 4096 candidates: %.1f ms = %.3g cand/s" % (1e3 * t, xs.shape[0] / t))
t0 = time.perf_counter(); [do.crit_SUR_cst(xi, fo, co, critcontrol=cc) for xi in xs[:3]]; tc = (time.perf_counter() - t0) / 3
print("oracle crit_SUR_cst (precomputation redone per point, as the reference does without critcontrol$precalc.data): %.1f ms per point" % (1e3 * tc))
