"""ctypes binding of the CPU oracle's binary32 build (oracle/_build/libkrrt_oracle_f32.so: krrt_oracle.c compiled with
every `double` a `float`, see the head of that file) — the checker of KRRT_F32 contexts.

Test infrastructure like oracle_binding.py.  The library's interface is float throughout; this module takes and returns
the binary64 arrays of the C ABI (inputs rounded to binary32 exactly as the kernels round them when they load, outputs
widened exactly) so that results compare byte for byte with libkrrt's."""
import ctypes
import os

import numpy as np

import oracle_binding as ob

ORACLE_F32_SO = os.path.join(ob.ORACLE_DIR, "_build", "libkrrt_oracle_f32.so")
F = ctypes.POINTER(ctypes.c_float)
I = ctypes.POINTER(ctypes.c_int32)


class KoConfigF(ctypes.Structure):
    _fields_ = [
        ("r", ctypes.c_float), ("R", ctypes.c_float), ("S", ctypes.c_int32), ("capacity", ctypes.c_int32),
        ("lo", ctypes.c_float * 4), ("hi", ctypes.c_float * 4),
        ("gamma", ctypes.c_float), ("rho_max", ctypes.c_float), ("p_direct", ctypes.c_float), ("goal_tol", ctypes.c_float),
        ("seed", ctypes.c_uint64), ("query_id", ctypes.c_uint32), ("use_kdtree", ctypes.c_int32), ("goal_mode", ctypes.c_int32),
    ]


RESULT_F32 = np.dtype([
    ("solved", "<i4"), ("goal_index", "<i4"), ("n_nodes", "<i4"), ("path_len", "<i4"),
    ("cost", "<f4"), ("time", "<f4"), ("waypoints", "<f4", (ob.KO_MAX_WAYPOINTS, 5))])

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    ob.build_oracle(force=not os.path.exists(ORACLE_F32_SO))
    L = ctypes.CDLL(ORACLE_F32_SO)
    L.ko_default_config.argtypes = [ctypes.POINTER(KoConfigF)]
    L.ko_steer_batch.argtypes = [F, F, ctypes.c_int, ctypes.c_float, F, F]
    L.ko_edge_collide.argtypes = [F, F, ctypes.c_float, ctypes.c_float, F, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                  ctypes.c_int]
    L.ko_edge_collide.restype = ctypes.c_int
    L.ko_solve.argtypes = [F, F, ctypes.c_float, ctypes.c_float, ctypes.c_float, F, ctypes.c_int]
    L.ko_solve.restype = ctypes.c_int
    L.ko_solve_inputs.argtypes = [F, F, ctypes.c_float, ctypes.c_float, F, ctypes.c_int]
    L.ko_solve_inputs.restype = ctypes.c_int
    L.ko_edge_state.argtypes = [F, F, ctypes.c_float, ctypes.c_float, F]
    L.ko_rho2.argtypes = [ctypes.POINTER(KoConfigF), ctypes.c_int]
    L.ko_rho2.restype = ctypes.c_float
    L.ko_sample.argtypes = [ctypes.POINTER(KoConfigF), ctypes.c_uint64, F, F]
    L.ko_plan_batch.argtypes = [ctypes.POINTER(KoConfigF), F, F, F, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                ctypes.c_uint32, ctypes.c_void_p, ctypes.c_int]
    L.ko_create.argtypes = [ctypes.POINTER(KoConfigF)]
    L.ko_create.restype = ctypes.c_void_p
    L.ko_destroy.argtypes = [ctypes.c_void_p]
    L.ko_set_obstacles.argtypes = [ctypes.c_void_p, F, ctypes.c_int, ctypes.c_int]
    L.ko_set_start.argtypes = [ctypes.c_void_p, F]
    L.ko_grow.argtypes = [ctypes.c_void_p, F, ctypes.c_int, F, I]
    L.ko_grow.restype = ctypes.c_int
    L.ko_size.argtypes = [ctypes.c_void_p]
    L.ko_size.restype = ctypes.c_int
    L.ko_iteration.argtypes = [ctypes.c_void_p]
    L.ko_iteration.restype = ctypes.c_int64
    for name, rt in (("ko_states", F), ("ko_ctc", F), ("ko_times", F), ("ko_last_sample", F), ("ko_parents", I)):
        getattr(L, name).argtypes = [ctypes.c_void_p]
        getattr(L, name).restype = rt
    L.ko_get_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(ob.KoStats)]
    L.ko_path.argtypes = [ctypes.c_void_p, ctypes.c_int, F, F, ctypes.c_int]
    L.ko_path.restype = ctypes.c_int
    L.ko_fill_result.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    L.ko_check_invariants.argtypes = [ctypes.c_void_p, ctypes.c_float]
    L.ko_check_invariants.restype = ctypes.c_int
    _lib = L
    return L


def _f(a):
    return a.ctypes.data_as(F)


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).astype(np.float32))


def default_config(**kw):
    c = KoConfigF()
    lib().ko_default_config(ctypes.byref(c))
    ob.set_fields(c, **kw)
    return c


def steer(x0, x1, r=0.5):
    """(tau*, c*) of x0[i] -> x1[i] in binary32, returned as binary64 arrays."""
    x0 = _f32(x0).reshape(-1, 4)
    x1 = _f32(x1).reshape(-1, 4)
    n = x0.shape[0]
    t = np.empty(n, dtype=np.float32)
    c = np.empty(n, dtype=np.float32)
    lib().ko_steer_batch(_f(x0), _f(x1), n, float(r), _f(t), _f(c))
    return t.astype(np.float64), c.astype(np.float64)


def edge_collide(x0, x1, tau, t0, obs, dynamic, R=1.35, S=10):
    x0 = _f32(x0)
    x1 = _f32(x1)
    obs = _f32(obs).reshape(-1, 4)
    return lib().ko_edge_collide(_f(x0), _f(x1), float(np.float32(tau)), float(np.float32(t0)), _f(obs), obs.shape[0],
                                 int(dynamic), float(R), int(S))


def solve(x0, x1, r=0.5, dt=0.1, t0=0.0, cap=4096):
    """rows (x, y, vx, vy, t0 + t) of solver.solve(x0, x1, dt) in binary32, widened"""
    x0, x1 = _f32(x0), _f32(x1)
    rows = np.empty((cap, 5), dtype=np.float32)
    n = lib().ko_solve(_f(x0), _f(x1), float(r), float(np.float32(dt)), float(np.float32(t0)), _f(rows), cap)
    return rows[:min(n, cap)].astype(np.float64)


def solve_inputs(x0, x1, r=0.5, dt=0.1, cap=4096):
    x0, x1 = _f32(x0), _f32(x1)
    rows = np.empty((cap, 2), dtype=np.float32)
    n = lib().ko_solve_inputs(_f(x0), _f(x1), float(r), float(np.float32(dt)), _f(rows), cap)
    return rows[:min(n, cap)].astype(np.float64)


def edge_state(x0, x1, tau, h):
    x0, x1 = _f32(x0), _f32(x1)
    out = np.empty(4, dtype=np.float32)
    lib().ko_edge_state(_f(x0), _f(x1), float(np.float32(tau)), float(np.float32(h)), _f(out))
    return out.astype(np.float64)


def rho2(cfg, n):
    return float(lib().ko_rho2(ctypes.byref(cfg), int(n)))


def sample(cfg, it, goal=None):
    out = np.empty(4, dtype=np.float32)
    g = _f32(goal) if goal is not None else None
    lib().ko_sample(ctypes.byref(cfg), int(it), _f(g) if g is not None else None, _f(out))
    return out.astype(np.float64)


def plan_batch(cfg, starts, goals, obstacles, dynamic, n_iter, query_base=0, nthreads=1):
    """ko_plan_batch of the float build; the records come back in the binary64 layout of krrt_result."""
    starts = _f32(starts).reshape(-1, 4)
    goals = _f32(goals).reshape(-1, 4)
    q = starts.shape[0]
    obstacles = _f32(obstacles)
    obstacles = obstacles.reshape(q, -1, 4) if obstacles.size else np.zeros((q, 0, 4), dtype=np.float32)
    m = obstacles.shape[1]
    raw = np.zeros(q, dtype=RESULT_F32)
    lib().ko_plan_batch(ctypes.byref(cfg), _f(starts), _f(goals), _f(obstacles), q, m, int(dynamic), int(n_iter),
                        int(query_base), raw.ctypes.data, int(nthreads))
    out = np.zeros(q, dtype=ob.RESULT_DTYPE)
    for k in ("solved", "goal_index", "n_nodes", "path_len"):
        out[k] = raw[k]
    for k in ("cost", "time", "waypoints"):
        out[k] = raw[k].astype(np.float64)
    return out


def _widen_result(raw):
    out = np.zeros(raw.shape, dtype=ob.RESULT_DTYPE)
    for k in ("solved", "goal_index", "n_nodes", "path_len"):
        out[k] = raw[k]
    for k in ("cost", "time", "waypoints"):
        out[k] = raw[k].astype(np.float64)
    return out


class OraclePlannerF32:
    """The single-tree planner of the float build (mirror of oracle_binding.OraclePlanner; arrays in and out are binary64)."""

    def __init__(self, **kw):
        self.cfg = default_config(**kw)
        self.h = lib().ko_create(ctypes.byref(self.cfg))

    def close(self):
        if self.h:
            lib().ko_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def set_obstacles(self, obs, dynamic):
        obs = _f32(obs).reshape(-1, 4)
        lib().ko_set_obstacles(self.h, _f(obs), obs.shape[0], int(dynamic))

    def set_start(self, x0):
        x0 = _f32(x0)
        lib().ko_set_start(self.h, _f(x0))

    def grow(self, goal, n_iter, samples=None):
        g = None
        if goal is not None:
            self._goal = _f32(goal)
            g = _f(self._goal)
        s = None
        if samples is not None:
            samples = _f32(samples).reshape(n_iter, 4)
            s = _f(samples)
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
        states = np.ctypeslib.as_array(L.ko_states(self.h), shape=(n, 4)).astype(np.float64)
        parent = np.ctypeslib.as_array(L.ko_parents(self.h), shape=(n,)).copy()
        ctc = np.ctypeslib.as_array(L.ko_ctc(self.h), shape=(n,)).astype(np.float64)
        tim = np.ctypeslib.as_array(L.ko_times(self.h), shape=(n,)).astype(np.float64)
        return states, parent, ctc, tim

    def last_sample(self):
        return np.ctypeslib.as_array(lib().ko_last_sample(self.h), shape=(4,)).astype(np.float64)

    def stats(self):
        s = ob.KoStats()
        lib().ko_get_stats(self.h, ctypes.byref(s))
        return {n: getattr(s, n) for n, _ in ob.KoStats._fields_}

    def near(self, x, radius2, cap=4096):
        x = _f32(x)
        idx = np.empty(cap, dtype=np.int32)
        nearest = ctypes.c_int32(-1)
        nd2 = ctypes.c_float(0)
        L = lib()
        L.ko_near.argtypes = [ctypes.c_void_p, F, ctypes.c_float, I, ctypes.c_int, ctypes.POINTER(ctypes.c_int32),
                              ctypes.POINTER(ctypes.c_float)]
        L.ko_near.restype = ctypes.c_int
        k = L.ko_near(self.h, _f(x), float(np.float32(radius2)), idx.ctypes.data_as(I), cap, ctypes.byref(nearest),
                      ctypes.byref(nd2))
        return idx[:min(k, cap)].copy(), nearest.value, float(nd2.value)

    def path(self, idx, cap=4096):
        st = np.empty((cap, 4), dtype=np.float32)
        tm = np.empty(cap, dtype=np.float32)
        n = lib().ko_path(self.h, int(idx), _f(st), _f(tm), cap)
        return st[:n].astype(np.float64), tm[:n].astype(np.float64)

    def result(self, solved, goal_index):
        r = np.zeros(1, dtype=RESULT_F32)
        lib().ko_fill_result(self.h, int(solved), int(goal_index), r.ctypes.data)
        return _widen_result(r)[0]

    def check_invariants(self, tol=1e-4):
        return lib().ko_check_invariants(self.h, float(tol))
