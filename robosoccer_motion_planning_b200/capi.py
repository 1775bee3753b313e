"""ctypes binding of libkrrt.so (include/krrt.h) — the Python host mirror of the C ABI.

There is no CPU fallback: if the shared library is missing, or no sm_100 device is present,
every constructor raises.  The names mirror the reference's component surface
(RRTStar::setStart / grow / goalIndex, Tree::states, solver.cost / solver.solve,
env.setObstacles — robosoccer_motion_planning_node.cpp:135-140,177,237-245).
"""
import ctypes
import os
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "lib", "libkrrt.so")

KRRT_MAX_OBS = 64
KRRT_MAX_WAYPOINTS = 32
KRRT_OK, KRRT_ERR_INVALID, KRRT_ERR_CUDA, KRRT_ERR_NO_DEVICE, KRRT_ERR_NEAR_OVERFLOW, KRRT_ERR_STATE, \
    KRRT_ERR_UNSUPPORTED = range(7)
KRRT_F64, KRRT_F32 = 0, 1   # krrt_config.precision
KRRT_STATIC_MAX_QUERIES, KRRT_STATIC_MAX_ITER = 65536, 1024   # limits of krrt_plan_batch_static

D = ctypes.POINTER(ctypes.c_double)
I = ctypes.POINTER(ctypes.c_int32)

# every symbol include/krrt.h declares (tests check the library exports all of them)
EXPORTS = (
    "krrt_default_config", "krrt_grow_done", "krrt_grow_logged", "krrt_create", "krrt_destroy", "krrt_last_error", "krrt_device_count",
    "krrt_set_obstacles", "krrt_set_start", "krrt_grow", "krrt_tree_size", "krrt_iteration", "krrt_last_sample",
    "krrt_get_stats", "krrt_tree_download", "krrt_path", "krrt_result_of", "krrt_trajectory", "krrt_steer_cost",
    "krrt_solve", "krrt_edge_collide", "krrt_near", "krrt_plan_batch", "krrt_plan_batch_device",
    "krrt_batch_stats", "krrt_launch_count", "krrt_set_profile", "krrt_last_timing", "krrt_timer_begin",
    "krrt_timer_end", "krrt_set_params", "krrt_tree_edges", "krrt_edge_samples", "krrt_kernel_times", "krrt_wave_stats", "krrt_tail_commits", "krrt_debug_clocks",
    "krrt_yaw_profiles", "krrt_yaw_profiles_device", "krrt_ingest_vision", "krrt_ingest_vision_device",
    "krrt_solve_inputs", "krrt_scan_stats", "krrt_tree_read", "krrt_goal_index",
    "krrt_plan_batch_device_chunks", "krrt_plan_batch_static", "krrt_host_alloc", "krrt_host_free", "krrt_flush_l2", "krrt_synchronize",
)


class KrrtConfig(ctypes.Structure):
    _fields_ = [
        ("device", ctypes.c_int32), ("precision", ctypes.c_int32), ("capacity", ctypes.c_int32),
        ("S", ctypes.c_int32), ("r", ctypes.c_double), ("R", ctypes.c_double),
        ("lo", ctypes.c_double * 4), ("hi", ctypes.c_double * 4),
        ("gamma", ctypes.c_double), ("rho_max", ctypes.c_double), ("p_direct", ctypes.c_double),
        ("goal_tol", ctypes.c_double), ("seed", ctypes.c_uint64), ("query_id", ctypes.c_uint32),
        ("batch", ctypes.c_int32), ("near_cap", ctypes.c_int32),
        ("goal_mode", ctypes.c_int32),
    ]


KRRT_GOAL_CONNECT, KRRT_GOAL_TOLERANCE = 0, 1


class KrrtStats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in (
        "iters", "inserted", "fail_reject", "fail_dup", "fail_nofree",
        "rewired", "propagated", "sum_near", "steer_calls", "collide_calls")]


RESULT_DTYPE = np.dtype([
    ("solved", "<i4"), ("goal_index", "<i4"), ("n_nodes", "<i4"), ("path_len", "<i4"),
    ("cost", "<f8"), ("time", "<f8"), ("waypoints", "<f8", (KRRT_MAX_WAYPOINTS, 5))])


ITER_RECORD_DTYPE = np.dtype([("n_nodes", "<i4"), ("solved", "<i4"), ("goal_index", "<i4"), ("inserted", "<i4"),
                              ("sample", "<f8", (4,)), ("goal_cost", "<f8")])   # == krrt_iter_record


class KrrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libkrrt error {code}: {msg}")
        self.code = code


def build_library(force=False):
    """nvcc-compile libkrrt.so for sm_100a in-tree (works without a GPU)."""
    srcs = [os.path.join(PKG_DIR, "csrc", f) for f in os.listdir(os.path.join(PKG_DIR, "csrc"))]
    srcs += [os.path.join(PKG_DIR, "Makefile"), os.path.join(PKG_DIR, "..", "include", "krrt.h")]
    if (not force and os.path.exists(LIB_PATH)
            and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    subprocess.check_call(["make", "-C", PKG_DIR, "-s", "all"])
    return LIB_PATH


_lib = None


def lib():
    """Load libkrrt.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KrrtError(-1, f"{LIB_PATH} is missing: run __graft_entry__.build() / make -C {PKG_DIR}; "
                            "there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    P = ctypes.c_void_p
    L.krrt_default_config.argtypes = [ctypes.POINTER(KrrtConfig)]
    L.krrt_default_config.restype = None
    L.krrt_create.argtypes = [ctypes.POINTER(KrrtConfig), ctypes.POINTER(P)]
    L.krrt_destroy.argtypes = [P]
    L.krrt_destroy.restype = None
    L.krrt_last_error.argtypes = [P]
    L.krrt_last_error.restype = ctypes.c_char_p
    L.krrt_device_count.argtypes = []
    L.krrt_set_obstacles.argtypes = [P, D, ctypes.c_int, ctypes.c_int]
    L.krrt_set_start.argtypes = [P, D]
    L.krrt_grow.argtypes = [P, D, ctypes.c_int, D, I, I]
    L.krrt_tree_size.argtypes = [P, I]
    L.krrt_grow_done.argtypes = [P, ctypes.POINTER(ctypes.c_int64), I]
    L.krrt_grow_logged.argtypes = [P, D, ctypes.c_int, D, ctypes.c_void_p, I, I]
    L.krrt_iteration.argtypes = [P, ctypes.POINTER(ctypes.c_int64)]
    L.krrt_last_sample.argtypes = [P, D]
    L.krrt_get_stats.argtypes = [P, ctypes.POINTER(KrrtStats)]
    L.krrt_batch_stats.argtypes = [P, ctypes.POINTER(KrrtStats)]
    L.krrt_tree_download.argtypes = [P, D, I, D, D]
    L.krrt_tree_read.argtypes = [P, ctypes.c_int, ctypes.c_int, D, I, D, D]
    L.krrt_goal_index.argtypes = [P, I]
    L.krrt_path.argtypes = [P, ctypes.c_int, D, D, ctypes.c_int, I]
    L.krrt_result_of.argtypes = [P, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    L.krrt_trajectory.argtypes = [P, ctypes.c_int, ctypes.c_double, D, ctypes.c_int, I]
    L.krrt_steer_cost.argtypes = [P, D, D, ctypes.c_int, D, D]
    L.krrt_solve.argtypes = [P, D, D, ctypes.c_double, ctypes.c_double, D, ctypes.c_int, I]
    L.krrt_solve_inputs.argtypes = [P, D, D, ctypes.c_double, ctypes.c_double, D, D, ctypes.c_int, I]
    L.krrt_edge_collide.argtypes = [P, D, D, D, D, ctypes.c_int, ctypes.POINTER(ctypes.c_ubyte)]
    L.krrt_near.argtypes = [P, D, ctypes.c_double, I, ctypes.c_int, I, I, D]
    L.krrt_plan_batch.argtypes = [P, D, D, D, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                  ctypes.c_uint32, ctypes.c_void_p]
    L.krrt_plan_batch_device.argtypes = [P, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint32,
                                         ctypes.c_void_p, ctypes.c_void_p]
    L.krrt_plan_batch_device_chunks.argtypes = [P, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                                ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint32,
                                                ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    L.krrt_plan_batch_static.argtypes = [ctypes.POINTER(KrrtConfig), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint32,
                                         ctypes.c_void_p, ctypes.c_void_p]
    L.krrt_host_alloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_uint64]
    L.krrt_host_free.argtypes = [ctypes.c_void_p]
    L.krrt_flush_l2.argtypes = [P]
    L.krrt_synchronize.argtypes = [P]
    L.krrt_set_params.argtypes = [P, ctypes.POINTER(KrrtConfig)]
    L.krrt_tree_edges.argtypes = [P, ctypes.c_int, D, ctypes.c_int]
    L.krrt_edge_samples.argtypes = [P, D, D, D, ctypes.c_int, ctypes.c_int, D]
    L.krrt_yaw_profiles.argtypes = [P, D, D, D, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int, D, D, D, D, I, I]
    L.krrt_yaw_profiles_device.argtypes = [P] + [ctypes.c_void_p] * 3 + [ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                                                         ctypes.c_int] + [ctypes.c_void_p] * 7
    F = ctypes.POINTER(ctypes.c_float)
    L.krrt_ingest_vision.argtypes = [P, F, I, ctypes.c_int, F, ctypes.c_int, I, ctypes.c_int, D, D, I, D, I]
    L.krrt_ingest_vision_device.argtypes = ([P, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 6)
    L.krrt_scan_stats.argtypes = [P, D, ctypes.POINTER(ctypes.c_int64), I]
    L.krrt_debug_clocks.argtypes = [P, ctypes.POINTER(ctypes.c_int64)]
    L.krrt_kernel_times.argtypes = [P, D, ctypes.POINTER(ctypes.c_int64)]
    L.krrt_wave_stats.argtypes = [P, ctypes.POINTER(ctypes.c_int64)]
    L.krrt_tail_commits.argtypes = [P, ctypes.POINTER(ctypes.c_int64)]
    L.krrt_launch_count.argtypes = [P]
    L.krrt_launch_count.restype = ctypes.c_int64
    L.krrt_set_profile.argtypes = [P, ctypes.c_int]
    L.krrt_timer_begin.argtypes = [P]
    L.krrt_timer_end.argtypes = [P, D]
    L.krrt_last_timing.argtypes = [P, D, D, ctypes.POINTER(ctypes.c_int64), D]
    _lib = L
    return L


def _d(a):
    return a.ctypes.data_as(D)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a if shape is None else a.reshape(shape)


class PinnedArray:
    """numpy view of page-locked host memory (krrt_host_alloc); keep the object alive as long as the array is used"""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * self.dtype.itemsize
        self.ptr = ctypes.c_void_p()
        rc = lib().krrt_host_alloc(ctypes.byref(self.ptr), max(n, 1))
        if rc != KRRT_OK:
            raise KrrtError(rc, "krrt_host_alloc failed")
        buf = (ctypes.c_ubyte * max(n, 1)).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=np.uint8, count=n).view(self.dtype).reshape(shape)

    @classmethod
    def like(cls, a):
        p = cls(a.shape, a.dtype)
        p.array[...] = a
        return p

    def __del__(self):
        try:
            if self.ptr:
                self.array = None
                lib().krrt_host_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


def default_config(**kw):
    c = KrrtConfig()
    lib().krrt_default_config(ctypes.byref(c))
    for k, v in kw.items():
        if k in ("lo", "hi"):
            for i in range(4):
                getattr(c, k)[i] = float(v[i])
        else:
            setattr(c, k, v)
    return c


def plan_batch_static(cfg, d_starts, d_goals, d_obstacles, q, m, dynamic, n_iter, query_base, d_out, stream=0):
    """krrt_plan_batch_static: the batched planner without a context and without a single allocation — for a caller
    inside a fixed-rate control cycle, or one whose device memory belongs to another allocator's pool (include/krrt.h).
    Device pointers as ints (tensor.data_ptr()); asynchronous on `stream` (0 = the default stream).  At most
    KRRT_STATIC_MAX_QUERIES queries of KRRT_STATIC_MAX_ITER iterations per call; calls share one scratch array per
    device, so do not let two of them overlap on the GPU.  Make the first call per device during start-up."""
    rc = lib().krrt_plan_batch_static(ctypes.byref(cfg), d_starts, d_goals, d_obstacles, int(q), int(m), int(dynamic),
                                      int(n_iter), int(query_base), d_out, stream or None)
    if rc != KRRT_OK:
        raise KrrtError(rc, lib().krrt_last_error(None).decode())


class Planner:
    """One CUDA-backed RRT* context (one device, one stream)."""

    def __init__(self, cfg=None, **kw):
        self.cfg = cfg if cfg is not None else default_config(**kw)
        if cfg is not None:
            for k, v in kw.items():
                setattr(self.cfg, k, v)
        self.h = ctypes.c_void_p()
        rc = lib().krrt_create(ctypes.byref(self.cfg), ctypes.byref(self.h))
        if rc != KRRT_OK:
            msg = lib().krrt_last_error(None).decode()
            self.h = None
            raise KrrtError(rc, msg)

    def _ck(self, rc):
        if rc != KRRT_OK:
            raise KrrtError(rc, lib().krrt_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            lib().krrt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # env.setObstacles (…node.cpp:139-141)
    def set_obstacles(self, obs, dynamic):
        obs = _f64(obs, (-1, 4))
        self._ck(lib().krrt_set_obstacles(self.h, _d(obs), obs.shape[0], int(dynamic)))

    # rrt.setStart + rrt.setIteration(0) (…node.cpp:135-136)
    def set_start(self, x0):
        x0 = _f64(x0, (4,))
        self._ck(lib().krrt_set_start(self.h, _d(x0)))

    # n_iter x rrt.grow(xg) (…node.cpp:176-177); returns (solved, goalIndex)
    def grow(self, goal, n_iter, samples=None):
        g = None
        if goal is not None:
            goal = _f64(goal, (4,))
            g = _d(goal)
        s = None
        if samples is not None:
            samples = _f64(samples, (n_iter, 4))
            s = _d(samples)
        solved = ctypes.c_int32(0)
        gi = ctypes.c_int32(-1)
        self._ck(lib().krrt_grow(self.h, g, int(n_iter), s, ctypes.byref(solved), ctypes.byref(gi)))
        return bool(solved.value), gi.value

    # the caller's `for(i<iteration) rrt.grow(xg)` loop as ONE device call (…node.cpp:176-185): per-iteration records
    def grow_logged(self, goal, n_iter, samples=None):
        g = None
        if goal is not None:
            goal = _f64(goal, (4,))
            g = _d(goal)
        s = None
        if samples is not None:
            samples = _f64(samples, (n_iter, 4))
            s = _d(samples)
        rec = np.zeros(n_iter, dtype=ITER_RECORD_DTYPE)
        solved = ctypes.c_int32(0)
        gi = ctypes.c_int32(-1)
        self._ck(lib().krrt_grow_logged(self.h, g, int(n_iter), s, rec.ctypes.data, ctypes.byref(solved), ctypes.byref(gi)))
        return rec, bool(solved.value), gi.value

    def grow_done(self):
        """(iterations the last grow call really ran, whether it stopped because the tree is full)"""
        it, full = ctypes.c_int64(0), ctypes.c_int32(0)
        self._ck(lib().krrt_grow_done(self.h, ctypes.byref(it), ctypes.byref(full)))
        return it.value, bool(full.value)

    @property
    def size(self):
        n = ctypes.c_int32(0)
        self._ck(lib().krrt_tree_size(self.h, ctypes.byref(n)))
        return n.value

    @property
    def iteration(self):
        n = ctypes.c_int64(0)
        self._ck(lib().krrt_iteration(self.h, ctypes.byref(n)))
        return n.value

    def last_sample(self):
        out = np.empty(4)
        self._ck(lib().krrt_last_sample(self.h, _d(out)))
        return out

    def stats(self):
        s = KrrtStats()
        self._ck(lib().krrt_get_stats(self.h, ctypes.byref(s)))
        return {n: getattr(s, n) for n, _ in KrrtStats._fields_}

    def batch_stats(self):
        s = KrrtStats()
        self._ck(lib().krrt_batch_stats(self.h, ctypes.byref(s)))
        return {n: getattr(s, n) for n, _ in KrrtStats._fields_}

    def tree(self):
        n = self.size
        states = np.empty((n, 4))
        parent = np.empty(n, dtype=np.int32)
        ctc = np.empty(n)
        tim = np.empty(n)
        self._ck(lib().krrt_tree_download(self.h, _d(states), parent.ctypes.data_as(I), _d(ctc), _d(tim)))
        return states, parent, ctc, tim

    def tree_read(self, first, count):
        """rows [first, first + count); may end with the connected goal node (index == size)"""
        states = np.empty((count, 4))
        parent = np.empty(count, dtype=np.int32)
        ctc = np.empty(count)
        tim = np.empty(count)
        self._ck(lib().krrt_tree_read(self.h, int(first), int(count), _d(states), parent.ctypes.data_as(I), _d(ctc),
                                      _d(tim)))
        return states, parent, ctc, tim

    # rrt.goalIndex() (robosoccer_visual.hpp:19)
    @property
    def goal_index(self):
        g = ctypes.c_int32(-1)
        self._ck(lib().krrt_goal_index(self.h, ctypes.byref(g)))
        return g.value

    # tree.states(idx) (…node.cpp:237)
    def path(self, idx, cap=4096):
        st = np.empty((cap, 4))
        tm = np.empty(cap)
        n = ctypes.c_int32(0)
        self._ck(lib().krrt_path(self.h, int(idx), _d(st), _d(tm), cap, ctypes.byref(n)))
        k = min(n.value, cap)
        return st[:k].copy(), tm[:k].copy()

    def result(self, solved, goal_index):
        r = np.zeros(1, dtype=RESULT_DTYPE)
        self._ck(lib().krrt_result_of(self.h, int(solved), int(goal_index), r.ctypes.data))
        return r[0]

    # get_trajectory(): every path edge re-solved at step dt (…node.cpp:232-249)
    def trajectory(self, idx, dt=0.1, cap_rows=1 << 16):
        rows = np.empty((cap_rows, 5))
        n = ctypes.c_int32(0)
        self._ck(lib().krrt_trajectory(self.h, int(idx), float(dt), _d(rows), cap_rows, ctypes.byref(n)))
        return rows[:min(n.value, cap_rows)].copy()

    # solver.cost(xi, xf) -> (t, c) (…node.cpp:242-244)
    def steer_cost(self, xi, xf):
        xi = _f64(xi, (-1, 4))
        xf = _f64(xf, (-1, 4))
        n = xi.shape[0]
        t = np.empty(n)
        c = np.empty(n)
        self._ck(lib().krrt_steer_cost(self.h, _d(xi), _d(xf), n, _d(t), _d(c)))
        return t, c

    # solver.solve(xi, xf, dt) (…node.cpp:245)
    def solve(self, xi, xf, dt=0.1, t0=0.0, cap_rows=4096):
        xi = _f64(xi, (4,))
        xf = _f64(xf, (4,))
        rows = np.empty((cap_rows, 5))
        n = ctypes.c_int32(0)
        self._ck(lib().krrt_solve(self.h, _d(xi), _d(xf), float(dt), float(t0), _d(rows), cap_rows, ctypes.byref(n)))
        return rows[:min(n.value, cap_rows)].copy()

    def solve_inputs(self, xi, xf, dt=0.1, t0=0.0, cap_rows=4096):
        """solve() plus the control inputs (ax, ay) at the same rows."""
        xi = _f64(xi, (4,))
        xf = _f64(xf, (4,))
        rows = np.empty((cap_rows, 5))
        inputs = np.empty((cap_rows, 2))
        n = ctypes.c_int32(0)
        self._ck(lib().krrt_solve_inputs(self.h, _d(xi), _d(xf), float(dt), float(t0), _d(rows), _d(inputs), cap_rows,
                                         ctypes.byref(n)))
        k = min(n.value, cap_rows)
        return rows[:k].copy(), inputs[:k].copy()

    # Connector slot: S samples (x,y,vx,vy,t0+t) of each optimal edge
    def edge_samples(self, xi, xf, t0, S=10):
        xi = _f64(xi, (-1, 4))
        xf = _f64(xf, (-1, 4))
        n = xi.shape[0]
        t0 = _f64(np.broadcast_to(t0, (n,)))
        rows = np.empty((n, S, 5))
        self._ck(lib().krrt_edge_samples(self.h, _d(xi), _d(xf), _d(t0), n, int(S), _d(rows)))
        return rows

    # tree.trajectories: the stored edge parent[i] -> i of every node
    def tree_edges(self, S=10):
        n = self.size
        rows = np.empty((n, S, 5))
        self._ck(lib().krrt_tree_edges(self.h, int(S), _d(rows), n))
        return rows

    # solver.set_weight / env.collision_radius / sampler settings on a live context
    def set_params(self, **kw):
        for k, v in kw.items():
            if k in ("lo", "hi"):
                for i in range(4):
                    getattr(self.cfg, k)[i] = float(v[i])
            else:
                setattr(self.cfg, k, v)
        self._ck(lib().krrt_set_params(self.h, ctypes.byref(self.cfg)))

    def edge_collide(self, xi, xf, tau, t0):
        xi = _f64(xi, (-1, 4))
        xf = _f64(xf, (-1, 4))
        n = xi.shape[0]
        tau = _f64(np.broadcast_to(tau, (n,)))
        t0 = _f64(np.broadcast_to(t0, (n,)))
        hit = np.zeros(n, dtype=np.uint8)
        self._ck(lib().krrt_edge_collide(self.h, _d(xi), _d(xf), _d(tau), _d(t0), n,
                                         hit.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte))))
        return hit

    def near(self, x, radius2, cap=4096):
        x = _f64(x, (4,))
        idx = np.empty(max(cap, 1), dtype=np.int32)
        cnt = ctypes.c_int32(0)
        nearest = ctypes.c_int32(-1)
        nd2 = ctypes.c_double(0)
        self._ck(lib().krrt_near(self.h, _d(x), float(radius2), idx.ctypes.data_as(I), cap, ctypes.byref(cnt),
                                 ctypes.byref(nearest), ctypes.byref(nd2)))
        return idx[:min(cnt.value, cap)].copy(), nearest.value, nd2.value

    # solve_rrt() for q independent queries (…node.cpp:167-196)
    def plan_batch(self, starts, goals, obstacles, dynamic, n_iter, query_base=0, out=None):
        """krrt_plan_batch on host buffers.  `out`: a caller-owned array of RESULT_DTYPE records (e.g. pinned memory, so
        the device-to-host copy of the 1304-byte records runs at full PCIe rate); allocated when None."""
        starts = _f64(starts, (-1, 4))
        goals = _f64(goals, (-1, 4))
        q = starts.shape[0]
        obstacles = np.asarray(obstacles, dtype=np.float64)
        obstacles = _f64(obstacles, (q, -1, 4)) if obstacles.size else np.zeros((q, 0, 4))
        m = obstacles.shape[1]
        if out is None:
            out = np.zeros(q, dtype=RESULT_DTYPE)
        assert out.dtype == RESULT_DTYPE and out.shape == (q,) and out.flags.c_contiguous
        self._ck(lib().krrt_plan_batch(self.h, _d(starts), _d(goals), _d(obstacles), q, m, int(dynamic), int(n_iter),
                                       int(query_base), out.ctypes.data))
        return out

    def plan_batch_device(self, d_starts, d_goals, d_obstacles, q, m, dynamic, n_iter, query_base, d_out, stream=0):
        """Device-pointer variant (ints from tensor.data_ptr()); asynchronous on `stream`."""
        self._ck(lib().krrt_plan_batch_device(self.h, d_starts, d_goals, d_obstacles, int(q), int(m), int(dynamic),
                                              int(n_iter), int(query_base), d_out, stream or None))

    def yaw_profiles(self, w, dw, yaw, v_max=1.0, a_max=1.0, n=25):
        """Trajectory1D::AngleController::angleControl + getAngleTrajectory (trajectory1d.cpp:326-346) for q queries."""
        w = np.ascontiguousarray(w, dtype=np.float64)
        dw = np.ascontiguousarray(dw, dtype=np.float64)
        yaw = np.ascontiguousarray(yaw, dtype=np.float64)
        q = w.shape[0]
        rows = max(int(n), 1)
        out = {"states": np.zeros((q, rows, 2)), "times": np.zeros((q, rows)), "inputs": np.zeros((q, rows)),
               "final_time": np.zeros(q), "status": np.zeros(q, dtype=np.int32), "n_seg": np.zeros(q, dtype=np.int32)}
        self._ck(lib().krrt_yaw_profiles(self.h, _d(w), _d(dw), _d(yaw), q, float(v_max), float(a_max), int(n),
                                         _d(out["states"]), _d(out["times"]), _d(out["inputs"]), _d(out["final_time"]),
                                         out["status"].ctypes.data_as(I), out["n_seg"].ctypes.data_as(I)))
        return out

    def yaw_profiles_device(self, d_w, d_dw, d_yaw, q, v_max, a_max, n, d_states, d_times, d_inputs, d_ft, d_status,
                            d_nseg, stream=0):
        self._ck(lib().krrt_yaw_profiles_device(self.h, d_w, d_dw, d_yaw, int(q), float(v_max), float(a_max), int(n),
                                                d_states, d_times, d_inputs, d_ft, d_status, d_nseg, stream or None))

    def ingest_vision(self, team, agent_id, rival, self_id):
        """RobotSubscriber::callback / rival_callback / getObstacles (robotsubscriber.cpp:29-77, robotsubscriber.h:50-55)
        for f frames: team f x n_team x 5 float32, agent_id f x n_team, rival f x n_rival x 4 float32, self_id f."""
        team = np.ascontiguousarray(team, dtype=np.float32)
        rival = np.ascontiguousarray(rival, dtype=np.float32)
        agent_id = np.ascontiguousarray(agent_id, dtype=np.int32)
        self_id = np.ascontiguousarray(self_id, dtype=np.int32)
        f = self_id.shape[0]
        n_team = team.shape[1] if team.ndim == 3 else 0
        n_rival = rival.shape[1] if rival.ndim == 3 else 0
        out = {"self_state": np.zeros((f, 4)), "heading": np.zeros(f), "self_index": np.zeros(f, dtype=np.int32),
               "obstacles": np.zeros((f, n_team + n_rival, 4)), "n_obs": np.zeros(f, dtype=np.int32)}
        F = ctypes.POINTER(ctypes.c_float)
        self._ck(lib().krrt_ingest_vision(self.h, team.ctypes.data_as(F), agent_id.ctypes.data_as(I), n_team,
                                          rival.ctypes.data_as(F), n_rival, self_id.ctypes.data_as(I), f,
                                          _d(out["self_state"]), _d(out["heading"]), out["self_index"].ctypes.data_as(I),
                                          _d(out["obstacles"]), out["n_obs"].ctypes.data_as(I)))
        return out

    def ingest_vision_device(self, d_team, d_agent_id, n_team, d_rival, n_rival, d_self_id, f, d_self_state, d_heading,
                             d_self_index, d_obstacles, d_n_obs, stream=0):
        self._ck(lib().krrt_ingest_vision_device(self.h, d_team, d_agent_id, int(n_team), d_rival, int(n_rival), d_self_id,
                                                 int(f), d_self_state, d_heading, d_self_index, d_obstacles, d_n_obs,
                                                 stream or None))

    def scan_stats(self):
        v, r, g = ctypes.c_double(0), ctypes.c_int64(0), ctypes.c_int32(0)
        self._ck(lib().krrt_scan_stats(self.h, ctypes.byref(v), ctypes.byref(r), ctypes.byref(g)))
        return {"visited_pairs": v.value, "index_rebuilds": r.value, "indexed_nodes": g.value}

    @property
    def launch_count(self):
        return lib().krrt_launch_count(self.h)

    def flush_l2(self):
        self._ck(lib().krrt_flush_l2(self.h))

    def synchronize(self):
        self._ck(lib().krrt_synchronize(self.h))

    def timer_begin(self):
        self._ck(lib().krrt_timer_begin(self.h))

    def timer_end(self):
        ms = ctypes.c_double(0)
        self._ck(lib().krrt_timer_end(self.h, ctypes.byref(ms)))
        return ms.value

    # phases of the persistent k_grow kernel (device clocks of its last CTA, scaled to the call's device time)
    KERNELS = ("scan", "steer", "decide", "validate", "apply", "sequential_windows", "-", "-")

    def kernel_times(self):
        ms = np.zeros(8)
        n = np.zeros(8, dtype=np.int64)
        self._ck(lib().krrt_kernel_times(self.h, _d(ms), n.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))))
        return {k: {"ms": float(ms[i]), "launches": int(n[i])} for i, k in enumerate(self.KERNELS)}

    def wave_stats(self):
        o = np.zeros(8, dtype=np.int64)
        self._ck(lib().krrt_wave_stats(self.h, o.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))))
        return {"rounds_parallel": int(o[0]), "rounds_sequential": int(o[1]), "expansions_parallel": int(o[2]),
                "prefix_ended_by": {"batch_end": int(o[3]), "in_batch_neighbour": int(o[4]), "sequential_sample": int(o[5]),
                                    "no_free_parent": int(o[6]), "rw_conflict": int(o[7])}}

    def tail_commits(self):
        o = ctypes.c_int64(0)
        self._ck(lib().krrt_tail_commits(self.h, ctypes.byref(o)))
        return int(o.value)

    def debug_clocks(self):
        o = np.zeros(16, dtype=np.int64)
        self._ck(lib().krrt_debug_clocks(self.h, o.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))))
        return o

    def set_profile(self, on):
        self._ck(lib().krrt_set_profile(self.h, int(on)))

    def last_timing(self):
        tot = ctypes.c_double(0)
        scan = ctypes.c_double(0)
        nl = ctypes.c_int64(0)
        work = ctypes.c_double(0)
        self._ck(lib().krrt_last_timing(self.h, ctypes.byref(tot), ctypes.byref(scan), ctypes.byref(nl),
                                        ctypes.byref(work)))
        return {"total_ms": tot.value, "scan_ms": scan.value, "scan_launches": nl.value,
                "scan_node_samples": work.value}
