"""CPU oracle for the DiceOptim acquisition hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; ``diceoptim_b200`` never does.

PARITY UNPINNED.  The reference (cran/DiceOptim 2.1.2, /root/reference) ships no tests, no
golden vectors and cannot be executed here (no R, no DiceKriging, no mnormt).  The arithmetic
of the path lives in the third-party package DiceKriging (``Depends: DiceKriging (>= 1.2)``,
/root/reference/DESCRIPTION:13; CRAN 1.6.0 assumed, source not vendored) and is restated in
``dicekriging.py`` from its published algorithm; ``diceoptim.py`` follows /root/reference/R/*.R
line by line.  The oracle is cross-checked by finite differences, Monte-Carlo and mpmath
(tests/test_oracle_*.py), which is the strongest pinning available without an R install.
"""
