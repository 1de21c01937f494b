import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def to_oracle(km):
    """diceoptim_b200.KM -> oracle.dicekriging.KM (same slots, independent classes)."""
    from oracle import dicekriging as dk
    c = km.covariance
    cov = dk.Covariance(name=c.name, sd2=c.sd2, range_val=np.asarray(c.range_val, float),
                        shape_val=None if c.shape_val is None else np.asarray(c.shape_val, float),
                        nugget=c.nugget, nugget_flag=c.nugget_flag)
    return dk.KM(X=km.X, y=km.y, F=km.F, trend=km.trend, trend_coef=km.trend_coef, covariance=cov,
                 noise_var=km.noise_var, T=km.T, z=km.z, M=km.M)


@pytest.fixture(scope="session")
def gpu_lib():
    import torch  # noqa: F401  (device plumbing only)
    from diceoptim_b200 import _lib
    return _lib.load()


def small_model(name="matern5_2", trend="const", n=60, d=3, seed=0, **kw):
    from diceoptim_b200 import synth
    X = synth.lhs(n, d, seed)
    y = np.sin(3.0 * X.sum(1)) + 0.5 * X[:, 0] ** 2
    theta = np.linspace(0.4, 0.7, d)
    shape = np.linspace(1.3, 1.9, d) if name == "powexp" else None
    sd2 = float(np.var(y)) if n > 2 else 1.0
    return synth.make_km(X, y, name, theta, sd2, trend=trend, shape=shape, **kw)
