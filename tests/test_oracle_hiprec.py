"""The float64 oracle against its extended-precision restatement (oracle/hiprec.py: long double + mpmath) on small
models: both follow the same reference lines, so they must agree to float64 rounding (amplified by the conditioning
of the factor).  This pins the adjudicator that tools/parity_table.py uses on the GPU box."""
import numpy as np
import pytest

from conftest import small_model, to_oracle


@pytest.mark.parametrize("name,trend,type", [("matern5_2", "const", "UK"), ("gauss", "linear", "UK"), ("exp", "const", "SK"),
                                             ("powexp", "linear", "SK"), ("matern3_2", "const", "UK")])
def test_hiprec_matches_oracle(name, trend, type):
    from oracle import dicekriging as dk, diceoptim as do, hiprec as hp
    km = to_oracle(small_model(name, trend, n=40, d=3, seed=3))
    x = np.random.default_rng(5).random((12, 3))
    ref = dk.predict_km(km, x, type=type)
    hi = hp.predict_km(km, x, type=type)
    assert np.max(np.abs(ref["mean"] - hi["mean"].astype(float))) < 1e-10
    assert np.max(np.abs(ref["sd"] - hi["sd"].astype(float)) / hi["sd"].astype(float)) < 1e-8
    ei = do.EI(x, km, type=type); ei_hi = hp.EI(x, km, type=type, pred=hi).astype(float)
    assert np.max(np.abs(ei - ei_hi)) <= 1e-8 * np.max(np.abs(ei_hi))
    g = do.EI_grad(x[3], km, type=type); g_hi = hp.EI_grad(x[3], km, type=type).astype(float)
    assert np.max(np.abs(g - g_hi)) <= 1e-7 * np.max(np.abs(g_hi))
    q = do.kriging_quantile_grad(x[4], km, beta=0.1, type=type)
    q_hi = hp.kriging_quantile_grad(x[4], km, do.qnorm(0.1), type=type).astype(float)
    assert np.max(np.abs(q - q_hi)) <= 1e-7 * np.max(np.abs(q_hi))


def test_hiprec_akg_matches_oracle():
    from oracle import diceoptim as do, hiprec as hp
    km = to_oracle(small_model("matern5_2", "const", n=30, d=2, seed=9, noise_var=0.05))
    x = np.random.default_rng(2).random((6, 2))
    predX = hp.predict_km(km, km.X, "UK")
    for xi in x:
        a = do.AKG(xi, km, new_noise_var=0.05); b = float(hp.AKG(xi, km, 0.05, predX=predX))
        assert abs(a - b) <= 1e-9 * max(abs(b), 1e-3)
