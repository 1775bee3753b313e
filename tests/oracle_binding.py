"""ctypes binding of the CPU oracle (oracle/krrt_oracle.c).

Test infrastructure: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (libkrrt.so) never touches it.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "libkrrt_oracle.so")

KO_MAX_WAYPOINTS = 32
D = ctypes.POINTER(ctypes.c_double)
I = ctypes.POINTER(ctypes.c_int32)


class KoConfig(ctypes.Structure):
    _fields_ = [
        ("r", ctypes.c_double),
        ("R", ctypes.c_double),
        ("S", ctypes.c_int32),
        ("capacity", ctypes.c_int32),
        ("lo", ctypes.c_double * 4),
        ("hi", ctypes.c_double * 4),
        ("gamma", ctypes.c_double),
        ("rho_max", ctypes.c_double),
        ("p_direct", ctypes.c_double),
        ("goal_tol", ctypes.c_double),
        ("seed", ctypes.c_uint64),
        ("query_id", ctypes.c_uint32),
        ("use_kdtree", ctypes.c_int32),
        ("goal_mode", ctypes.c_int32),
    ]


class KoResult(ctypes.Structure):
    _fields_ = [
        ("solved", ctypes.c_int32),
        ("goal_index", ctypes.c_int32),
        ("n_nodes", ctypes.c_int32),
        ("path_len", ctypes.c_int32),
        ("cost", ctypes.c_double),
        ("time", ctypes.c_double),
        ("waypoints", (ctypes.c_double * 5) * KO_MAX_WAYPOINTS),
    ]


class KoStats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in (
        "iters", "inserted", "fail_reject", "fail_dup", "fail_nofree",
        "rewired", "propagated", "sum_near", "steer_calls", "collide_calls")]


RESULT_DTYPE = np.dtype([
    ("solved", "<i4"), ("goal_index", "<i4"), ("n_nodes", "<i4"), ("path_len", "<i4"),
    ("cost", "<f8"), ("time", "<f8"), ("waypoints", "<f8", (KO_MAX_WAYPOINTS, 5))])
assert RESULT_DTYPE.itemsize == ctypes.sizeof(KoResult)


def build_oracle(force=False):
    src = [os.path.join(ORACLE_DIR, f) for f in ("krrt_oracle.c", "krrt_oracle.h", "Makefile")]
    if (not force and os.path.exists(ORACLE_SO)
            and all(os.path.getmtime(ORACLE_SO) >= os.path.getmtime(s) for s in src)):
        return ORACLE_SO
    subprocess.check_call(["make", "-C", ORACLE_DIR, "-s", "all"])
    return ORACLE_SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = ctypes.CDLL(build_oracle())
    L.ko_default_config.argtypes = [ctypes.POINTER(KoConfig)]
    L.ko_steer_batch.argtypes = [D, D, ctypes.c_int, ctypes.c_double, D, D]
    L.ko_edge_state.argtypes = [D, D, ctypes.c_double, ctypes.c_double, D]
    L.ko_solve.argtypes = [D, D, ctypes.c_double, ctypes.c_double, ctypes.c_double, D, ctypes.c_int]
    L.ko_solve.restype = ctypes.c_int
    L.ko_set_near_radius.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_double]
    L.ko_set_near_radius.restype = None
    L.ko_solve_inputs.argtypes = [D, D, ctypes.c_double, ctypes.c_double, D, ctypes.c_int]
    L.ko_solve_inputs.restype = ctypes.c_int
    L.ko_edge_collide.argtypes = [D, D, ctypes.c_double, ctypes.c_double, D, ctypes.c_int, ctypes.c_int,
                                  ctypes.c_double, ctypes.c_int]
    L.ko_edge_collide.restype = ctypes.c_int
    L.ko_sample.argtypes = [ctypes.POINTER(KoConfig), ctypes.c_uint64, D, D]
    L.ko_philox.argtypes = [ctypes.c_uint64] + [ctypes.c_uint32] * 4 + [ctypes.POINTER(ctypes.c_uint32)]
    L.ko_rho2.argtypes = [ctypes.POINTER(KoConfig), ctypes.c_int]
    L.ko_rho2.restype = ctypes.c_double
    L.ko_create.argtypes = [ctypes.POINTER(KoConfig)]
    L.ko_create.restype = ctypes.c_void_p
    L.ko_destroy.argtypes = [ctypes.c_void_p]
    L.ko_set_obstacles.argtypes = [ctypes.c_void_p, D, ctypes.c_int, ctypes.c_int]
    L.ko_set_start.argtypes = [ctypes.c_void_p, D]
    L.ko_grow.argtypes = [ctypes.c_void_p, D, ctypes.c_int, D, I]
    L.ko_grow.restype = ctypes.c_int
    L.ko_size.argtypes = [ctypes.c_void_p]
    L.ko_size.restype = ctypes.c_int
    L.ko_iteration.argtypes = [ctypes.c_void_p]
    L.ko_iteration.restype = ctypes.c_int64
    for name, rt in (("ko_states", D), ("ko_ctc", D), ("ko_times", D), ("ko_last_sample", D), ("ko_parents", I)):
        getattr(L, name).argtypes = [ctypes.c_void_p]
        getattr(L, name).restype = rt
    L.ko_get_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(KoStats)]
    L.ko_near.argtypes = [ctypes.c_void_p, D, ctypes.c_double, I, ctypes.c_int, I, D]
    L.ko_near.restype = ctypes.c_int
    L.ko_path.argtypes = [ctypes.c_void_p, ctypes.c_int, D, D, ctypes.c_int]
    L.ko_path.restype = ctypes.c_int
    L.ko_fill_result.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(KoResult)]
    L.ko_check_invariants.argtypes = [ctypes.c_void_p, ctypes.c_double]
    L.ko_check_invariants.restype = ctypes.c_int
    L.ko_plan_batch.argtypes = [ctypes.POINTER(KoConfig), D, D, D, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                ctypes.c_int, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_int]
    _lib = L
    return L


def _d(a):
    return a.ctypes.data_as(D)


def default_config(**kw):
    c = KoConfig()
    lib().ko_default_config(ctypes.byref(c))
    set_fields(c, **kw)
    return c


def set_fields(c, **kw):
    for k, v in kw.items():
        if k in ("lo", "hi"):
            for i in range(4):
                getattr(c, k)[i] = float(v[i])
        else:
            setattr(c, k, v)
    return c


def steer(x0, x1, r=0.5):
    x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(-1, 4)
    x1 = np.ascontiguousarray(x1, dtype=np.float64).reshape(-1, 4)
    n = x0.shape[0]
    tau = np.empty(n)
    cost = np.empty(n)
    lib().ko_steer_batch(_d(x0), _d(x1), n, float(r), _d(tau), _d(cost))
    return tau, cost


def edge_state(x0, x1, tau, h):
    """state on the optimal edge at h = t/tau; scalars -> (4,), arrays (n,4),(n,4),(n,),(n,) -> (n,4)"""
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    x1 = np.ascontiguousarray(x1, dtype=np.float64)
    if x0.ndim == 1:
        out = np.empty(4)
        lib().ko_edge_state(_d(x0), _d(x1), float(tau), float(h), _d(out))
        return out
    n = x0.shape[0]
    tau = np.broadcast_to(tau, (n,))
    h = np.broadcast_to(h, (n,))
    out = np.empty((n, 4))
    f = lib().ko_edge_state
    for i in range(n):
        f(_d(x0[i]), _d(x1[i]), float(tau[i]), float(h[i]), _d(out[i]))
    return out


def solve(x0, x1, r=0.5, dt=0.1, t0=0.0, cap=4096):
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    x1 = np.ascontiguousarray(x1, dtype=np.float64)
    rows = np.empty((cap, 5))
    n = lib().ko_solve(_d(x0), _d(x1), float(r), float(dt), float(t0), _d(rows), cap)
    return rows[:min(n, cap)].copy()


def edge_collide(x0, x1, tau, t0, obs, dynamic, R=1.35, S=10):
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    x1 = np.ascontiguousarray(x1, dtype=np.float64)
    obs = np.ascontiguousarray(obs, dtype=np.float64).reshape(-1, 4)
    return lib().ko_edge_collide(_d(x0), _d(x1), float(tau), float(t0), _d(obs), obs.shape[0], int(dynamic),
                                 float(R), int(S))


def sample(cfg, it, goal=None):
    out = np.empty(4)
    g = None if goal is None else _d(np.ascontiguousarray(goal, dtype=np.float64))
    lib().ko_sample(ctypes.byref(cfg), int(it), g, _d(out))
    return out


class OraclePlanner:
    """Mirrors the C-ABI planner (krrt.h) on the CPU."""

    def __init__(self, cfg=None, **kw):
        self.cfg = cfg if cfg is not None else default_config(**kw)
        if cfg is not None:
            set_fields(self.cfg, **kw)
        self.h = lib().ko_create(ctypes.byref(self.cfg))

    def close(self):
        if self.h:
            lib().ko_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def set_obstacles(self, obs, dynamic):
        obs = np.ascontiguousarray(obs, dtype=np.float64).reshape(-1, 4)
        lib().ko_set_obstacles(self.h, _d(obs), obs.shape[0], int(dynamic))

    def set_params(self, gamma, rho_max):
        lib().ko_set_near_radius(self.h, float(gamma), float(rho_max))

    def set_start(self, x0):
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        lib().ko_set_start(self.h, _d(x0))

    def grow(self, goal, n_iter, samples=None):
        g = None
        if goal is not None:
            self._goal = np.ascontiguousarray(goal, dtype=np.float64)
            g = _d(self._goal)
        s = None
        if samples is not None:
            samples = np.ascontiguousarray(samples, dtype=np.float64).reshape(n_iter, 4)
            s = _d(samples)
        gi = ctypes.c_int32(-1)
        solved = lib().ko_grow(self.h, g, int(n_iter), s, ctypes.byref(gi))
        return bool(solved), gi.value

    @property
    def size(self):
        return lib().ko_size(self.h)

    @property
    def iteration(self):
        return lib().ko_iteration(self.h)

    def tree(self):
        n = self.size
        L = lib()
        states = np.ctypeslib.as_array(L.ko_states(self.h), shape=(n, 4)).copy()
        parent = np.ctypeslib.as_array(L.ko_parents(self.h), shape=(n,)).copy()
        ctc = np.ctypeslib.as_array(L.ko_ctc(self.h), shape=(n,)).copy()
        tim = np.ctypeslib.as_array(L.ko_times(self.h), shape=(n,)).copy()
        return states, parent, ctc, tim

    def last_sample(self):
        return np.ctypeslib.as_array(lib().ko_last_sample(self.h), shape=(4,)).copy()

    def stats(self):
        s = KoStats()
        lib().ko_get_stats(self.h, ctypes.byref(s))
        return {n: getattr(s, n) for n, _ in KoStats._fields_}

    def near(self, x, radius2, cap=4096):
        x = np.ascontiguousarray(x, dtype=np.float64)
        idx = np.empty(cap, dtype=np.int32)
        nearest = ctypes.c_int32(-1)
        nd2 = ctypes.c_double(0)
        k = lib().ko_near(self.h, _d(x), float(radius2), idx.ctypes.data_as(I), cap, ctypes.byref(nearest),
                          ctypes.byref(nd2))
        return idx[:min(k, cap)].copy(), nearest.value, nd2.value

    def path(self, idx, cap=4096):
        st = np.empty((cap, 4))
        tm = np.empty(cap)
        n = lib().ko_path(self.h, int(idx), _d(st), _d(tm), cap)
        return st[:n].copy(), tm[:n].copy()

    def result(self, solved, goal_index):
        r = np.zeros(1, dtype=RESULT_DTYPE)
        lib().ko_fill_result(self.h, int(solved), int(goal_index), r.ctypes.data_as(ctypes.POINTER(KoResult)))
        return r[0]

    def check_invariants(self, tol=1e-9):
        return lib().ko_check_invariants(self.h, float(tol))


def plan_batch(cfg, starts, goals, obstacles, dynamic, n_iter, query_base=0, nthreads=1):
    starts = np.ascontiguousarray(starts, dtype=np.float64).reshape(-1, 4)
    goals = np.ascontiguousarray(goals, dtype=np.float64).reshape(-1, 4)
    q = starts.shape[0]
    obstacles = np.ascontiguousarray(obstacles, dtype=np.float64)
    obstacles = obstacles.reshape(q, -1, 4) if obstacles.size else np.zeros((q, 0, 4))
    m = obstacles.shape[1]
    out = np.zeros(q, dtype=RESULT_DTYPE)
    lib().ko_plan_batch(ctypes.byref(cfg), _d(starts), _d(goals), _d(obstacles), q, m, int(dynamic), int(n_iter),
                        int(query_base), out.ctypes.data, int(nthreads))
    return out
