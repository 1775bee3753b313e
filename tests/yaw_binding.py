"""ctypes bindings of the yaw-profile checkers: oracle/yaw_oracle.c (restatement) and, when it has been built,
oracle/_ref/libyaw_ref.so (the REFERENCE's own trajectory1d.cpp).  Test infrastructure only."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
YAW_SO = os.path.join(ROOT, "oracle", "_build", "libyaw_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libyaw_ref.so")
D = ctypes.POINTER(ctypes.c_double)
I = ctypes.POINTER(ctypes.c_int32)

_ARGS = [D, D, D, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int, D, D, D, D, I, I]


def _d(a):
    return a.ctypes.data_as(D)


def _load(path, sym, sp):
    L = ctypes.CDLL(path)
    f = getattr(L, sym)
    f.argtypes = _ARGS
    f.restype = None
    g = getattr(L, sp)
    g.argtypes = [ctypes.c_double, ctypes.c_double, D, D]
    g.restype = ctypes.c_double
    return f, g


def have_ref():
    return os.path.exists(REF_SO)


def _profiles(fn, w, dw, yaw, v_max, a_max, n):
    w = np.ascontiguousarray(w, dtype=np.float64)
    dw = np.ascontiguousarray(dw, dtype=np.float64)
    yaw = np.ascontiguousarray(yaw, dtype=np.float64)
    q = w.shape[0]
    rows = max(n, 1)
    out = {"states": np.zeros((q, rows, 2)), "times": np.zeros((q, rows)), "inputs": np.zeros((q, rows)),
           "final_time": np.zeros(q), "status": np.zeros(q, dtype=np.int32), "n_seg": np.zeros(q, dtype=np.int32)}
    fn(_d(w), _d(dw), _d(yaw), q, float(v_max), float(a_max), int(n), _d(out["states"]), _d(out["times"]),
       _d(out["inputs"]), _d(out["final_time"]), out["status"].ctypes.data_as(I), out["n_seg"].ctypes.data_as(I))
    return out


def oracle_profiles(w, dw, yaw, v_max=1.0, a_max=1.0, n=25):
    f, _ = _load(YAW_SO, "yo_yaw_profiles", "yo_shortest_path")
    return _profiles(f, w, dw, yaw, v_max, a_max, n)


def ref_profiles(w, dw, yaw, v_max=1.0, a_max=1.0, n=25):
    f, _ = _load(REF_SO, "yr_yaw_profiles", "yr_shortest_path")
    return _profiles(f, w, dw, yaw, v_max, a_max, n)


def shortest_path(which, a, b):
    _, g = _load(YAW_SO if which == "oracle" else REF_SO, "yo_shortest_path" if which == "oracle" else "yr_shortest_path",
                 "yo_shortest_path" if which == "oracle" else "yr_shortest_path")
    fo, to = ctypes.c_double(0), ctypes.c_double(0)
    d = g(float(a), float(b), ctypes.byref(fo), ctypes.byref(to))
    return d, fo.value, to.value


def random_queries(rng, q, w_range=4.0, dw_range=2.5):
    """Initial heading / rate and target yaw as the planner sees them: angles a little beyond [-pi, pi], rates up to and
    beyond v_max (case 3), a share of queries at rest or already on target."""
    w = rng.uniform(-w_range, w_range, q)
    dw = rng.uniform(-dw_range, dw_range, q)
    yaw = rng.uniform(-w_range, w_range, q)
    dw[::5] = 0.0
    yaw[::17] = w[::17]
    dw[::23] = 1.0   # exactly v_max: the cruising case
    return w, dw, yaw
