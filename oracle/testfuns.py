"""DiceOptim's analytic test functions on [0,1]^d (used only to synthesise y for fixtures/bench)."""
import numpy as np


def branin2(x):
    """R/branin2.R:36-45."""
    x = np.atleast_2d(x)
    m, s = 54.8104, 51.9496
    xx = 15.0 * x[:, 0] - 5.0; y = 15.0 * x[:, 1]
    f = (y - 5.1 * xx ** 2 / (4 * np.pi ** 2) + 5 * xx / np.pi - 6) ** 2 + 10 * (1 - 1 / (8 * np.pi)) * np.cos(xx) + 10
    return (f - m) / s


def hartman4(x):
    """R/hartman4.R:33-58 (rows 1-4 of the 6-dim Hartman tables, first 4 columns)."""
    x = np.atleast_2d(x)
    a = np.array([[10.0, 3.0, 17.0, 3.5], [0.05, 10.0, 17.0, 0.1], [3.0, 3.5, 1.7, 10.0], [17.0, 8.0, 0.05, 10.0]])
    p = np.array([[0.1312, 0.1696, 0.5569, 0.0124], [0.2329, 0.4135, 0.8307, 0.3736],
                  [0.2348, 0.1451, 0.3522, 0.2883], [0.4047, 0.8828, 0.8732, 0.5743]])
    c = np.array([1.0, 1.2, 3.0, 3.2])
    d = ((x[:, None, :] - p[None]) ** 2 * a[None]).sum(-1)
    return (-(c[None] * np.exp(-d)).sum(-1) + 1.1) / 0.8387


def rosenbrock4(x):
    """R/rosenbrock4.R:36-45."""
    x = np.atleast_2d(x)
    m, s = 382658.057227524, 375264.858362295
    xx = 15.0 * x - 5.0
    f = (100.0 * (xx[:, :-1] ** 2 - xx[:, 1:]) ** 2 + (1 - xx[:, :-1]) ** 2).sum(1)
    return (f - m) / s


def sphere_w(x):
    """'sphere6-style' weighted shifted sphere generalised to any d:
    sum_i 2^((i mod 6)+1) (10 x_i - 5 + s_i)^2 with fixed shifts, standardised empirically."""
    x = np.atleast_2d(x)
    d = x.shape[1]
    i = np.arange(d)
    w = 2.0 ** ((i % 6) + 1)
    s = 0.5 * np.cos(1.0 + i)
    f = (w * (10.0 * x - 5.0 + s) ** 2).sum(1)
    mu = (w * (100.0 / 12.0 + s ** 2)).sum()
    return (f - mu) / mu
