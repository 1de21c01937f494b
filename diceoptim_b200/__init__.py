"""diceoptim_b200 — B200-native batched acquisition engine behind DiceOptim's function API.

The package holds only the hot path of SURVEY.md section 8: csrc/ (CUDA kernels + C ABI,
include/dk_b200.h), the ctypes binding, and the host-side mirror of the R interface
(EI, EI_grad, kriging_quantile[_grad], screen_EI / max_crit, ...).  Model fitting stays in
DiceKriging; ``synth`` only builds km objects with FIXED hyper-parameters for tests and bench.
"""
from .km import KM, Covariance, DeviceModel  # noqa: F401
from .api import (AEI, AEI_grad, EQI_grad, AKG, AKG_grad, EI, EQI, EI_grad, kriging_quantile, kriging_quantile_grad, max_crit, max_EI, max_qEI, predict, qEI, qEI_grad,  # noqa: F401
                  sampleFromEI, screen_EI, checkPredict, crit_EFI, crit_SUR_cst, Feas, prob_of_feasible_improvement,
                  max_AEI, max_AKG, max_EQI, min_quantile, integration_design_cst)

__all__ = ["AEI", "AEI_grad", "EQI", "EQI_grad", "AKG", "AKG_grad", "KM", "Covariance", "DeviceModel", "EI", "EI_grad", "kriging_quantile", "kriging_quantile_grad",
           "max_crit", "max_EI", "max_qEI", "predict", "qEI", "qEI_grad", "sampleFromEI", "screen_EI", "checkPredict", "crit_EFI", "crit_SUR_cst", "Feas",
           "prob_of_feasible_improvement", "max_AEI", "max_AKG", "max_EQI", "min_quantile", "integration_design_cst"]
