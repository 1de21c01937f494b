"""Golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from the oracle; the reference has
none).  CPU: the oracle still reproduces them bit-for-bit-ish (1e-12), which pins it against silent
edits.  GPU: the CUDA path matches them at the parity tolerances."""
import os

import numpy as np
import pytest

from conftest import to_oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    from diceoptim_b200 import KM, Covariance
    z = np.load(os.path.join(G, name))
    nv = z["noise_var"]
    cov = Covariance(name=str(z["kernel"]), sd2=float(z["sd2"]), range_val=z["theta"], nugget=float(z["nugget"]),
                     nugget_flag=bool(z["nugget_flag"]))
    km = KM(X=z["X"], y=z["y"], F=z["F"], trend=str(z["trend"]), trend_coef=z["beta"], covariance=cov, T=z["T"], z=z["z"],
            M=z["M"], noise_var=None if nv.size == 0 else nv)
    return z, km


def close(a, b, rtol, atol):
    a = np.asarray(a); b = np.asarray(b)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(b)
    assert np.all(np.abs(a[ok] - b[ok]) <= rtol * np.abs(b[ok]) + atol), np.max(np.abs(a[ok] - b[ok]))


def test_oracle_reproduces_golden():
    from oracle import dicekriging as dk, diceoptim as do
    z, km = load("ei_branin_fact9.npz"); ko = to_oracle(km)
    with np.errstate(all="ignore"):
        close(do.EI(z["x"], ko), z["EI"], 1e-12, 1e-15)
    close(dk.predict_km(ko, z["x"])["sd"], z["sd"], 1e-12, 1e-15)
    close(do.EI_grad(z["x"][37], ko), z["EI_grad"][37], 1e-12, 1e-15)
    z, km = load("akg_branin_noisy12.npz"); ko = to_oracle(km)
    close([do.AKG(x, ko, new_noise_var=float(z["new_noise_var"])) for x in z["x"][::13]], z["AKG"][::13], 1e-12, 1e-15)
    z, km = load("qei_hartman4_81.npz"); ko = to_oracle(km)
    close([do.qEI(b, ko, plugin=float(z["plugin"]), fastCompute=False) for b in z["xb4"][:3]], z["qEI4"][:3], 1e-12, 1e-15)


@pytest.mark.gpu
def test_cuda_matches_golden(gpu_lib):
    from diceoptim_b200 import AKG, EI, EI_grad, kriging_quantile, kriging_quantile_grad, predict, qEI
    z, km = load("ei_branin_fact9.npz")
    sc = np.max(np.abs(km.y)) + np.sqrt(km.covariance.sd2)
    pr = predict(km, z["x"])
    close(pr["mean"], z["mean"], 0, 1e-10 * sc)
    close(pr["sd"] ** 2, z["sd"] ** 2, 0, 1e-10 * km.covariance.sd2)
    close(EI(z["x"], km), z["EI"], 1e-9, 1e-12 * np.nanmax(z["EI"]))
    close(EI_grad(z["x"], km), z["EI_grad"], 1e-8, 1e-11 * np.nanmax(np.abs(z["EI_grad"])))
    close(kriging_quantile(z["x"], km, beta=0.1), z["q10"], 0, 1e-10 * sc)
    close(kriging_quantile_grad(z["x"], km, beta=0.1), z["q10_grad"], 1e-8, 1e-9 * np.max(np.abs(z["q10_grad"])))
    km.release()
    z, km = load("akg_branin_noisy12.npz")
    close(AKG(z["x"], km, new_noise_var=float(z["new_noise_var"])), z["AKG"], 1e-8, 1e-11)
    km.release()
    z, km = load("qei_hartman4_81.npz")
    for q in (2, 3, 4):
        close(qEI(z[f"xb{q}"], km, plugin=float(z["plugin"])), z[f"qEI{q}"], 1e-9, 1e-12)
    km.release()


def load_cst():
    from diceoptim_b200 import KM, Covariance
    z = np.load(os.path.join(G, "cst_branin_16.npz"))
    models = []
    for tag in ("o", "c1", "c2"):
        cov = Covariance(name=str(z[f"{tag}_kernel"]), sd2=float(z[f"{tag}_sd2"]), range_val=z[f"{tag}_theta"],
                         nugget=float(z[f"{tag}_nugget"]), nugget_flag=bool(z[f"{tag}_nugget_flag"]))
        models.append(KM(X=z[f"{tag}_X"], y=z[f"{tag}_y"], F=z[f"{tag}_F"], trend=str(z[f"{tag}_trend"]),
                         trend_coef=z[f"{tag}_beta"], covariance=cov, T=z[f"{tag}_T"], z=z[f"{tag}_z"], M=z[f"{tag}_M"],
                         noise_var=z[f"{tag}_noise_var"]))
    return z, models[0], models[1:]


def test_oracle_reproduces_constrained_golden():
    from oracle import diceoptim as do
    z, fun, cst = load_cst()
    fo, co = to_oracle(fun), [to_oracle(m) for m in cst]
    idx = [0, 17, 40, 62, 80]
    close(do.check_predict_distance(z["x"][idx], co + [fo], "covdist"), z["mindist"][idx], 1e-12, 1e-15)
    close([do.crit_EFI(z["x"][i], fo, co) for i in idx], z["EFI"][idx], 1e-12, 1e-18)
    close(do.prob_of_feasible_improvement(z["x"][idx], fo, co), z["PFI"][idx], 1e-12, 1e-18)
    close([do.crit_SUR_cst(z["x"][i], fo, co, critcontrol={"integration.points": z["P"]}) for i in idx[:3]], z["SUR"][idx[:3]],
          1e-12, 1e-20)


@pytest.mark.gpu
def test_cuda_matches_constrained_golden(gpu_lib):
    from diceoptim_b200 import checkPredict, crit_EFI, crit_SUR_cst, prob_of_feasible_improvement
    z, fun, cst = load_cst()
    close(checkPredict(z["x"], cst + [fun], return_distance=True), z["mindist"], 1e-9, 1e-9)
    close(crit_EFI(z["x"], fun, cst), z["EFI"], 1e-9, 1e-14)
    close(prob_of_feasible_improvement(z["x"], fun, cst), z["PFI"], 1e-9, 1e-14)
    close(crit_SUR_cst(z["x"], fun, cst, critcontrol={"integration.points": z["P"]}), z["SUR"], 1e-9, 1e-15)
    for m in cst + [fun]:
        m.release()
