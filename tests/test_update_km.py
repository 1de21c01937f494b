"""update_km (n^2 append of one observation) against a from-scratch rebuild of T, z, M with the same covariance
parameters and trend coefficients - what computeAuxVariables does in the reference's constant-liar loop
(R/max_qEI.R:305-316).  1e-12 absolute on factors of order 1.  CPU only."""
import numpy as np
import pytest

from diceoptim_b200 import synth
from diceoptim_b200.update import update_km


@pytest.mark.parametrize("trend", ["const", "linear"])
@pytest.mark.parametrize("kw", [{}, {"nugget": 1e-3}, {"noise_var": 0.02}])
def test_append_matches_rebuild(trend, kw):
    X = synth.lhs(40, 3, 1); y = np.sin(X.sum(1))
    km = synth.make_km(X, y, "matern5_2", 0.5, 1.0, trend=trend, **kw)
    cur = km
    pts = np.random.default_rng(0).random((3, 3))
    for i, xn in enumerate(pts):
        cur = update_km(cur, xn, 0.7 + i)
    nv = None if km.noise_var is None else np.append(km.noise_var, np.zeros(3))
    ref = synth.make_km(np.vstack([X, pts]), np.append(y, [0.7, 1.7, 2.7]), "matern5_2", 0.5, 1.0, trend=trend,
                        beta=km.trend_coef, nugget=kw.get("nugget"), noise_var=nv)
    assert cur.n == 43 and np.array_equal(cur.trend_coef, km.trend_coef)
    for a, b in ((cur.T, ref.T), (cur.z, ref.z), (cur.M, ref.M), (cur.F, ref.F), (cur.X, ref.X), (cur.y, ref.y)):
        assert np.max(np.abs(a - b)) < 1e-12
    assert np.allclose(np.tril(cur.T, -1), 0.0)


def test_duplicate_point_is_refused():
    X = synth.lhs(20, 2, 2); y = X.sum(1)
    km = synth.make_km(X, y, "gauss", 0.4, 1.0)
    with pytest.raises(np.linalg.LinAlgError):
        update_km(km, X[3], 0.0)
    with pytest.raises(ValueError):
        update_km(km, np.zeros(3), 0.0)
