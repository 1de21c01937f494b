"""One call of checkPredict (covdist) and of crit_SUR_cst on mid-size inputs, for an ncu pass over k_covdist / k_sur_*.
usage: python tools/profile_constrained.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from diceoptim_b200 import checkPredict, crit_SUR_cst, synth

n, d = 2000, 10
X = synth.lhs(n, d, 4)
yo = synth.sphere_w(X); yo = (yo - yo.mean()) / yo.std()
yc = X[:, 0] + X[:, 1] - 1.0 + 0.2 * np.cos(5 * X[:, 2])
nv = np.full(n, 1e-3)
fun = synth.make_km(X, yo, "matern5_2", np.full(d, 1.0), 1.0, noise_var=nv)
c1 = synth.make_km(X, yc, "matern5_2", np.full(d, 1.0), float(np.var(yc)), noise_var=nv)
rng = np.random.default_rng(0)
x = rng.random((296 * 64, d))
for _ in range(int(os.environ.get("PROFILE_REPS", "2"))):
    checkPredict(x, [c1, fun])
    crit_SUR_cst(x[:4096], fun, [c1], critcontrol={"integration.points": rng.random((2048, d))})
print("profile_constrained done")
