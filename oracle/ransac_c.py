"""ctypes binding of oracle/ransac_c.c.  TEST INFRASTRUCTURE / timed CPU baseline only."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_ransac.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "ransac_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_ransac.restype = ctypes.c_int
        _lib.orc_fit_all.restype = ctypes.c_int
        _lib.orc_num_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads():
    return lib().orc_num_threads()


def fit_all(src, tgt, minimal_sets):
    src = np.ascontiguousarray(src, np.float64)
    tgt = np.ascontiguousarray(tgt, np.float64)
    ms = np.ascontiguousarray(minimal_sets, np.int32)
    H, n = ms.shape
    T = np.empty((H, 4, 4), np.float64)
    lib().orc_fit_all(_p(src), _p(tgt), _p(ms), ctypes.c_int(H), ctypes.c_int(n), _p(T))
    return T


def ransac(src, tgt, max_dist, minimal_sets, adaptive=False, confidence=0.999, return_all=False):
    src = np.ascontiguousarray(src, np.float64)
    tgt = np.ascontiguousarray(tgt, np.float64)
    ms = np.ascontiguousarray(minimal_sets, np.int32)
    H, n = ms.shape
    T = np.empty((4, 4), np.float64)
    best_h, inl = ctypes.c_int32(-1), ctypes.c_int32(0)
    s = ctypes.c_double(0.0)
    counts = np.zeros(H, np.int32) if return_all else None
    sums = np.zeros(H, np.float64) if return_all else None
    n_eval = lib().orc_ransac(_p(src), _p(tgt), ctypes.c_int(len(src)), _p(ms), ctypes.c_int(H),
                              ctypes.c_int(n), ctypes.c_double(max_dist), ctypes.c_int(int(adaptive)),
                              ctypes.c_double(confidence), _p(T), ctypes.byref(best_h),
                              ctypes.byref(inl), ctypes.byref(s),
                              _p(counts) if return_all else None, _p(sums) if return_all else None)
    out = dict(T=T, best_h=best_h.value, inliers=inl.value, sum_d2=s.value, n_evaluated=n_eval)
    if return_all:
        out["counts"], out["sums"] = counts[:n_eval], sums[:n_eval]
    return out
