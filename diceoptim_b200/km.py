"""Host-side mirror of the DiceKriging ``km`` object: the S4 slots DiceOptim's hot path reads
(read sites /root/reference/R/EI.grad.R:118-122, R/AKG.R:110-113,124,
R/qEI.grad.R:78-83) plus the device handle that dk_model_create() returns for them."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _lib

KERNEL_IDS = {"gauss": 0, "exp": 1, "matern3_2": 2, "matern5_2": 3, "powexp": 4}
TREND_IDS = {"const": 0, "linear": 1}
TYPE_IDS = {"SK": 0, "UK": 1}


@dataclass
class Covariance:
    """covTensorProduct: name, sd2, range.val, shape.val, nugget, nugget.flag."""
    name: str
    sd2: float
    range_val: np.ndarray
    shape_val: Optional[np.ndarray] = None
    nugget: float = 0.0
    nugget_flag: bool = False


@dataclass
class KM:
    """km slots: X, y, F, trend.formula ('const' = ~1, 'linear' = ~.), trend.coef, covariance,
    noise.var, T (upper Cholesky), z, M."""
    X: np.ndarray
    y: np.ndarray
    F: np.ndarray
    trend: str
    trend_coef: np.ndarray
    covariance: Covariance
    T: np.ndarray
    z: np.ndarray
    M: np.ndarray
    noise_var: Optional[np.ndarray] = None
    _handles: dict = field(default_factory=dict, repr=False, compare=False)

    @property
    def n(self): return self.X.shape[0]
    @property
    def d(self): return self.X.shape[1]
    @property
    def p(self): return self.F.shape[1]

    def handle(self, device: int = 0) -> "DeviceModel":
        """Device-resident copy of the model on CUDA device ``device`` (created once, cached)."""
        h = self._handles.get(device)
        if h is None:
            h = DeviceModel(self, device)
            self._handles[device] = h
        return h

    def release(self):
        for h in self._handles.values():
            h.close()
        self._handles.clear()


def _f64(a, order="F"):
    return np.require(np.asarray(a, dtype=np.float64), requirements=["ALIGNED"] + (["F"] if order == "F" else ["C"]))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_lib.c_double_p)


def make_desc(km: "KM", device: int = 0):
    """(dk_model_desc, keep-alive dict of the arrays it points to) for a km object."""
    if km.covariance.name not in KERNEL_IDS:
        raise ValueError(f"covariance kernel {km.covariance.name!r} is not a covTensorProduct kernel "
                         f"the engine supports {sorted(KERNEL_IDS)}")
    if km.trend not in TREND_IDS:
        raise ValueError(f"trend {km.trend!r} not supported (const = ~1, linear = ~.)")
    cov = km.covariance
    keep = dict(
        theta=_f64(np.broadcast_to(cov.range_val, (km.d,)).copy()),
        shape=None if cov.shape_val is None else _f64(cov.shape_val),
        noise_var=None if km.noise_var is None else _f64(km.noise_var),
        T=_f64(km.T), z=_f64(km.z), M=_f64(np.atleast_2d(km.M).reshape(km.n, km.p)),
        X=_f64(km.X), y=_f64(km.y), F=_f64(km.F), beta=_f64(km.trend_coef),
    )
    desc = _lib.dk_model_desc(
        n=km.n, d=km.d, p=km.p, kernel=KERNEL_IDS[cov.name], trend=TREND_IDS[km.trend], sd2=float(cov.sd2),
        theta=_ptr(keep["theta"]), shape=_ptr(keep["shape"]), nugget=float(cov.nugget),
        nugget_flag=int(bool(cov.nugget_flag)), noise_var=_ptr(keep["noise_var"]),
        T=_ptr(keep["T"]), z=_ptr(keep["z"]), M=_ptr(keep["M"]), X=_ptr(keep["X"]), y=_ptr(keep["y"]),
        F=_ptr(keep["F"]), beta=_ptr(keep["beta"]), device=device)
    return desc, keep


class MultiDeviceModel:
    """Owns one dk_multi* (include/dk_b200.h): the model replicated on several GPUs of this process.
    ``devices``: list of CUDA ordinals, or an int = the first that many visible devices (0: all)."""

    def __init__(self, km: "KM", devices=0):
        lib = _lib.load()
        desc, keep = make_desc(km)
        out = C.c_void_p()
        if isinstance(devices, int):
            rc = lib.dk_multi_create(C.byref(desc), None, devices, C.byref(out))
        else:
            arr = (C.c_int * len(devices))(*devices)
            rc = lib.dk_multi_create(C.byref(desc), arr, len(devices), C.byref(out))
        _lib.check(rc, "dk_multi_create")
        self._h, self._lib = out, lib
        self.n, self.d, self.p = km.n, km.d, km.p

    @property
    def ptr(self):
        if self._h is None:
            raise _lib.DkError("multi-device model already destroyed")
        return self._h

    def count(self) -> int:
        return int(self._lib.dk_multi_count(self.ptr))

    def close(self):
        if self._h is not None:
            self._lib.dk_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceModel:
    """Owns one dk_model* (include/dk_b200.h).  Not thread-safe; calls on it must not overlap."""

    def __init__(self, km: KM, device: int = 0):
        lib = _lib.load()
        self.n, self.d, self.p, self.device = km.n, km.d, km.p, device
        desc, keep = make_desc(km, device)
        out = C.c_void_p()
        _lib.check(lib.dk_model_create(C.byref(desc), C.byref(out)), "dk_model_create")
        self._h = out
        self._lib = lib

    @property
    def ptr(self):
        if self._h is None:
            raise _lib.DkError("device model already destroyed")
        return self._h

    def clone(self, device=None) -> "DeviceModel":
        """A replica of this handle (device-to-device copy of the tiled model, no re-tiling on the host)."""
        new = object.__new__(DeviceModel)
        out = C.c_void_p()
        dev = self.device if device is None else int(device)
        _lib.check(self._lib.dk_model_clone(self.ptr, dev, C.byref(out)), "dk_model_clone")
        new._h, new._lib = out, self._lib
        new.n, new.d, new.p, new.device = self.n, self.d, self.p, dev
        return new

    def device_bytes(self) -> int:
        return int(self._lib.dk_model_device_bytes(self.ptr))

    def launches(self) -> int:
        return int(self._lib.dk_model_launches(self.ptr))

    def close(self):
        if self._h is not None:
            self._lib.dk_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
