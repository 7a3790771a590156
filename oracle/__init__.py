"""ctypes front-end of oracle/glm_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  photon_ml_b200 never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libglm_oracle.so")

LOGISTIC, POISSON, SQUARED, SMOOTHED_HINGE = 0, 1, 2, 3


def build(force=False):
    """Compile glm_oracle.c with gcc (a few hundred ms)."""
    src = os.path.join(_HERE, "glm_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libglm_oracle.so"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_log1p_exp.restype = ctypes.c_double
        _lib.oracle_log1p_exp.argtypes = [ctypes.c_double]
        _lib.oracle_dzz.restype = ctypes.c_double
        _lib.oracle_dzz.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double]
        _lib.oracle_loss_and_dz.restype = None
        _lib.oracle_loss_and_dz.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                            ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Data:
    """Rows in CSR, values widened to fp64 exactly as the reference holds them."""

    def __init__(self, n, d, row_ptr, col, val, label, offset=None, weight=None):
        self.n, self.d = int(n), int(d)
        self.row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int64)
        self.col = np.ascontiguousarray(col, dtype=np.int32)
        self.val = np.ascontiguousarray(val, dtype=np.float64)
        self.label = np.ascontiguousarray(label, dtype=np.float64)
        self.offset = None if offset is None else np.ascontiguousarray(offset, dtype=np.float64)
        self.weight = None if weight is None else np.ascontiguousarray(weight, dtype=np.float64)
        assert self.row_ptr.shape == (self.n + 1,)
        assert self.label.shape == (self.n,)

    def _args(self):
        return [ctypes.c_int64(self.n), ctypes.c_int32(self.d), _p(self.row_ptr), _p(self.col), _p(self.val),
                _p(self.label), _p(self.offset), _p(self.weight)]


def _norm(d, factors, shifts, intercept):
    f = None if factors is None else np.ascontiguousarray(factors, dtype=np.float64)
    s = None if shifts is None else np.ascontiguousarray(shifts, dtype=np.float64)
    if f is not None and f.shape != (d,):
        raise ValueError("factors size mismatch")
    if s is not None and s.shape != (d,):
        raise ValueError("shifts size mismatch")
    return f, s, ctypes.c_int32(-1 if intercept is None else int(intercept))


def log1p_exp(x):
    return lib().oracle_log1p_exp(float(x))


def loss_and_dz(loss, margin, label):
    l, dl = ctypes.c_double(), ctypes.c_double()
    lib().oracle_loss_and_dz(loss, float(margin), float(label), ctypes.byref(l), ctypes.byref(dl))
    return l.value, dl.value


def dzz(loss, margin, label):
    return lib().oracle_dzz(loss, float(margin), float(label))


def value_and_gradient(data, loss, coef, factors=None, shifts=None, intercept=None, parts=None, arbiter=False):
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    if coef.shape != (data.d,):
        raise ValueError("coefficient size mismatch")
    f, s, icpt = _norm(data.d, factors, shifts, intercept)
    value = ctypes.c_double()
    grad = np.empty(data.d, dtype=np.float64)
    args = data._args() + [ctypes.c_int(loss), _p(coef), _p(f), _p(s), icpt]
    if parts is not None:
        rc = lib().oracle_value_and_gradient_partitioned(*args, ctypes.c_int(int(parts)), ctypes.byref(value),
                                                         _p(grad))
    elif arbiter:
        rc = lib().oracle_value_and_gradient_arbiter(*args, ctypes.byref(value), _p(grad))
    else:
        rc = lib().oracle_value_and_gradient(*args, ctypes.byref(value), _p(grad))
    if rc:
        raise ValueError("requirement failed (intercept factor/shift or shift without intercept)")
    return value.value, grad


def hessian_vector(data, loss, coef, v, factors=None, shifts=None, intercept=None):
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    v = np.ascontiguousarray(v, dtype=np.float64)
    if coef.shape != (data.d,) or v.shape != (data.d,):
        raise ValueError("size mismatch")
    f, s, icpt = _norm(data.d, factors, shifts, intercept)
    hv = np.empty(data.d, dtype=np.float64)
    rc = lib().oracle_hessian_vector(*data._args(), ctypes.c_int(loss), _p(coef), _p(v), _p(f), _p(s), icpt, _p(hv))
    if rc:
        raise ValueError("requirement failed" if rc == 1 else "loss is not twice differentiable")
    return hv


def hessian_diagonal(data, loss, coef):
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    if coef.shape != (data.d,):
        raise ValueError("size mismatch")
    diag = np.empty(data.d, dtype=np.float64)
    rc = lib().oracle_hessian_diagonal(*data._args(), ctypes.c_int(loss), _p(coef), _p(diag))
    if rc:
        raise ValueError("loss is not twice differentiable")
    return diag


def max_threads():
    return lib().oracle_max_threads()
