"""Generates tests/golden/*.npz from the ORACLE (the reference itself cannot run here: no R, no DiceKriging;
SURVEY 8(c) "fixtures to author").  The set-ups mirror the reference's own examples with FIXED
hyper-parameters (the examples fit them by MLE, which is DiceKriging's job and out of scope):
  ei_branin_fact9   : 3 x 3 factorial design on branin, gauss, theta = (0.5, 0.5), 12 x 12 grid   (R/EI.R:55-73)
  akg_branin_noisy12: 12-point LHS, noisy branin2, noise.var = 0.2, 12 x 12 grid                 (R/AKG.R:27-47)
  qei_hartman4_81   : 3^4 factorial design on hartman4, matern5_2, q = 2, 3, 4 batches             (R/qEI.R:71-104)
  cst_branin_16     : 4 x 4 design, branin objective + two noisy constraint models; checkPredict distance, crit_EFI,
                      prob.of.feasible.improvement, crit_SUR_cst on a 9 x 9 grid          (R/crit_EFI.R:60-118)
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import dicekriging as dk, diceoptim as do, testfuns  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def model_arrays(m):
    c = m.covariance
    return dict(X=m.X, y=m.y, F=m.F, T=m.T, z=m.z, M=m.M, beta=m.trend_coef, theta=c.range_val, sd2=c.sd2,
                nugget=c.nugget, nugget_flag=c.nugget_flag, noise_var=np.array([]) if m.noise_var is None else m.noise_var)


def grid(g, d=2):
    u = np.linspace(0, 1, g)
    return np.array(np.meshgrid(*([u] * d), indexing="ij")).reshape(d, -1).T[:, ::-1].copy()   # first coord fastest


def main():
    # 1. EI example
    g3 = np.array([0.0, 0.5, 1.0])
    X = np.array([[a, b] for b in g3 for a in g3])
    y = testfuns.branin2(X)
    m = dk.make_km(X, y, "gauss", [0.5, 0.5], float(np.var(y, ddof=1)))
    xg = grid(12)
    pr = dk.predict_km(m, xg)
    with np.errstate(all="ignore"):
        ei = do.EI(xg, m)
    gr = np.array([do.EI_grad(x, m) for x in xg])
    q10 = do.kriging_quantile(xg, m, beta=0.1)
    gq = np.array([do.kriging_quantile_grad(x, m, beta=0.1) for x in xg])
    np.savez(os.path.join(OUT, "ei_branin_fact9.npz"), kernel="gauss", trend="const", x=xg, mean=pr["mean"], sd=pr["sd"],
             EI=ei, EI_grad=gr, q10=q10, q10_grad=gq, **model_arrays(m))
    # 2. AKG example
    rng = np.random.default_rng(421)
    X = (np.argsort(rng.random((12, 2)), axis=0) + rng.random((12, 2))) / 12
    y = testfuns.branin2(X) + np.sqrt(0.2) * rng.standard_normal(12)
    m = dk.make_km(X, y, "matern5_2", [0.4, 0.6], float(np.var(y, ddof=1)), noise_var=0.2)
    akg = np.array([do.AKG(x, m, new_noise_var=0.2) for x in xg])
    np.savez(os.path.join(OUT, "akg_branin_noisy12.npz"), kernel="matern5_2", trend="const", x=xg, AKG=akg,
             new_noise_var=0.2, **model_arrays(m))
    # 3. qEI example
    X = np.array(np.meshgrid(g3, g3, g3, g3, indexing="ij")).reshape(4, -1).T
    y = testfuns.hartman4(X)
    m = dk.make_km(X, y, "matern5_2", [0.6] * 4, float(np.var(y, ddof=1)))
    rng = np.random.default_rng(7)
    plugin = float(np.median(y))
    out = {}
    for q in (2, 3, 4):
        xb = rng.random((12, q, 4))
        out[f"xb{q}"] = xb
        out[f"qEI{q}"] = np.array([do.qEI(b, m, plugin=plugin, fastCompute=False) for b in xb])
    np.savez(os.path.join(OUT, "qei_hartman4_81.npz"), kernel="matern5_2", trend="const", plugin=plugin, **out, **model_arrays(m))
    # 4. constrained example: objective + two constraints, all with noise.var (covdist then has a non-trivial kn_xy)
    g4 = np.linspace(0.0, 1.0, 4)
    X = np.array([[a, b] for b in g4 for a in g4])
    yo = testfuns.branin2(X)
    yc1 = X[:, 0] * X[:, 1] - 0.3
    yc2 = 0.15 - (X[:, 0] - 0.5) ** 2 - 0.5 * X[:, 1]
    nv = np.full(16, 0.01)
    mo = dk.make_km(X, yo, "matern5_2", [0.5, 0.5], float(np.var(yo, ddof=1)), noise_var=nv)
    m1 = dk.make_km(X, yc1, "gauss", [0.6, 0.6], float(np.var(yc1, ddof=1)), noise_var=nv)
    m2 = dk.make_km(X, yc2, "matern3_2", [0.7, 0.5], float(np.var(yc2, ddof=1)), trend="linear", noise_var=nv)
    xg9 = grid(9)
    P = np.random.default_rng(99).random((30, 2))
    cc = {"integration.points": P}
    out = dict(x=xg9, P=P,
               mindist=do.check_predict_distance(xg9, [m1, m2, mo], "covdist"),
               EFI=np.array([do.crit_EFI(x, mo, [m1, m2]) for x in xg9]),
               PFI=do.prob_of_feasible_improvement(xg9, mo, [m1, m2]),
               SUR=np.array([do.crit_SUR_cst(x, mo, [m1, m2], critcontrol=cc) for x in xg9]))
    for tag, m, kern, tr in (("o", mo, "matern5_2", "const"), ("c1", m1, "gauss", "const"), ("c2", m2, "matern3_2", "linear")):
        out.update({f"{tag}_{k}": v for k, v in model_arrays(m).items()})
        out[f"{tag}_kernel"] = kern; out[f"{tag}_trend"] = tr
    np.savez(os.path.join(OUT, "cst_branin_16.npz"), **out)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
