"""Monte-Carlo sweep of independent planning queries over the GPUs of one box (BASELINE config 5):
ctypes mirror of libkrrt_mgpu.so (include/krrt_mgpu.h).  No torch, no CPU fallback.

Queries are independent (own start / goal / obstacles / Philox query id), so the path shards with NO data-path
collective: GPU g plans the contiguous block [g*Q/G, (g+1)*Q/G) and the only exchange is the gather of the fixed-size
result records (krrt_result, 1304 B each) over NCCL, overlapped with the planning chunk by chunk inside the library.
Results do not depend on the number of GPUs because the Philox stream is keyed by the GLOBAL query id.

Two ways to stand it up (krrt_mgpu.h): `Sweep.single_process(n_gpus)` drives all GPUs from this process;
`Sweep.from_env()` is one rank of a `python -m torch.distributed.run` / mpirun style launch (RANK, WORLD_SIZE,
LOCAL_RANK, MASTER_PORT in the environment): rank 0 makes the NCCL id and hands it over through a file.
"""
import ctypes
import os
import time

import numpy as np

from . import capi
from .capi import RESULT_DTYPE, KrrtConfig, KrrtError, KrrtStats, KRRT_OK

MGPU_LIB_PATH = os.path.join(capi.PKG_DIR, "lib", "libkrrt_mgpu.so")
ID_BYTES = 128

# every symbol include/krrt_mgpu.h declares (tests check the library exports all of them)
EXPORTS = (
    "krrt_mgpu_create", "krrt_mgpu_unique_id", "krrt_mgpu_create_rank", "krrt_mgpu_destroy", "krrt_mgpu_last_error",
    "krrt_mgpu_info", "krrt_mgpu_shard", "krrt_mgpu_chunk_layout", "krrt_sweep_multi_gpu", "krrt_sweep_load",
    "krrt_sweep_run", "krrt_sweep_results", "krrt_sweep_stats", "krrt_mgpu_launch_count", "krrt_sweep_overlap",
    "krrt_mgpu_barrier", "krrt_mgpu_max", "krrt_sweep_last_timing",
)

_lib = None


def lib():
    """Load libkrrt_mgpu.so (and with it libkrrt.so and NCCL); raises when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    capi.lib()
    if not os.path.exists(MGPU_LIB_PATH):
        raise KrrtError(-1, f"{MGPU_LIB_PATH} is missing: run __graft_entry__.build() / make -C {capi.PKG_DIR}")
    L = ctypes.CDLL(MGPU_LIB_PATH)
    P, D, I = ctypes.c_void_p, capi.D, capi.I
    L.krrt_mgpu_create.argtypes = [ctypes.POINTER(KrrtConfig), ctypes.c_int, I, ctypes.POINTER(P)]
    L.krrt_mgpu_unique_id.argtypes = [ctypes.c_void_p]
    L.krrt_mgpu_create_rank.argtypes = [ctypes.POINTER(KrrtConfig), ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.POINTER(P)]
    L.krrt_mgpu_destroy.argtypes = [P]
    L.krrt_mgpu_destroy.restype = None
    L.krrt_mgpu_last_error.argtypes = [P]
    L.krrt_mgpu_last_error.restype = ctypes.c_char_p
    L.krrt_mgpu_info.argtypes = [P, I, I, I, I, I]
    L.krrt_mgpu_shard.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, I, I]
    L.krrt_mgpu_shard.restype = None
    L.krrt_mgpu_chunk_layout.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                         ctypes.POINTER(ctypes.c_int64), I, I]
    L.krrt_sweep_multi_gpu.argtypes = [P, D, D, D, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_uint32, ctypes.c_void_p]
    L.krrt_sweep_load.argtypes = [P, D, D, D, ctypes.c_int, ctypes.c_int]
    L.krrt_sweep_run.argtypes = [P, ctypes.c_int, ctypes.c_int, ctypes.c_uint32, D]
    L.krrt_sweep_results.argtypes = [P, ctypes.c_void_p]
    L.krrt_sweep_stats.argtypes = [P, ctypes.POINTER(KrrtStats)]
    L.krrt_mgpu_launch_count.argtypes = [P]
    L.krrt_mgpu_launch_count.restype = ctypes.c_int64
    L.krrt_sweep_overlap.argtypes = [P, I, I]
    L.krrt_sweep_last_timing.argtypes = [P, D, D]
    L.krrt_mgpu_barrier.argtypes = [P]
    L.krrt_mgpu_max.argtypes = [P, D]
    _lib = L
    return L


def shard_bounds(q_total, rank, world):
    """Contiguous block of rank `rank` (sizes differ by at most one query): krrt_mgpu_shard."""
    lo, hi = ctypes.c_int32(0), ctypes.c_int32(0)
    lib().krrt_mgpu_shard(int(q_total), int(rank), int(world), ctypes.byref(lo), ctypes.byref(hi))
    return lo.value, hi.value


def chunk_layout(q_total, world, chunk, rank, c):
    """(offset, count, global index) of chunk c of rank `rank` in the gathered [chunk][rank][records] buffer."""
    off, cnt, glob = ctypes.c_int64(0), ctypes.c_int32(0), ctypes.c_int32(0)
    rc = lib().krrt_mgpu_chunk_layout(int(q_total), int(world), int(chunk), int(rank), int(c), ctypes.byref(off),
                                      ctypes.byref(cnt), ctypes.byref(glob))
    if rc != KRRT_OK:
        raise KrrtError(rc, "krrt_mgpu_chunk_layout: bad arguments")
    return off.value, cnt.value, glob.value


def unpermute(gathered, q_total, world, chunk):
    """Host restatement of what krrt_sweep_results does with the gathered buffer: records in global query order."""
    pad = (q_total + world - 1) // world
    n_chunks = max(1, (pad + chunk - 1) // chunk)
    out = np.zeros(q_total, dtype=gathered.dtype)
    for r in range(world):
        for c in range(n_chunks):
            off, cnt, glob = chunk_layout(q_total, world, chunk, r, c)
            out[glob:glob + cnt] = gathered[off:off + cnt]
    return out


def _launcher():
    """(pid, start time in s since the epoch) of the process that started this rank — torchrun's agent, mpirun, a test.
    All ranks of one launch share it and two launches do not, which MASTER_PORT and TORCHELASTIC_RUN_ID cannot say
    (torchrun's defaults are 29500 and 'none' every time).  Start time None where /proc cannot tell."""
    ppid = os.getppid()
    try:
        with open(f"/proc/{ppid}/stat") as f:
            ticks = int(f.read().rsplit(")", 1)[1].split()[19])       # field 22: starttime in clock ticks since boot
        with open("/proc/stat") as f:
            btime = next(int(line.split()[1]) for line in f if line.startswith("btime"))
        return ppid, btime + ticks / os.sysconf("SC_CLK_TCK")
    except (OSError, ValueError, StopIteration, IndexError):
        return ppid, None


def _id_path(tag=None):
    ppid, born = _launcher()
    tag = tag or os.environ.get("KRRT_MGPU_JOB") or "{}_{}_{}_{}".format(
        os.environ.get("MASTER_PORT", "0"), os.environ.get("TORCHELASTIC_RUN_ID", "x"), ppid,
        int(born) if born is not None else "x")
    return os.path.join(os.environ.get("KRRT_ID_DIR", "/tmp"), f"krrt_mgpu_id_{tag}.bin"), born


def exchange_id(rank, world, tag=None, timeout=120.0, make_id=None):
    """Rank 0 makes the NCCL id and leaves it in a file every rank of the box can read (one node: /tmp).

    The file belongs to ONE job: its name carries the launcher's pid and start time (or KRRT_MGPU_JOB, a nonce the
    launcher exports), rank 0 removes whatever is there under that name before it asks NCCL for a new id, readers take
    only a file written after their launcher was started, and release_id() removes it once the communicator is up."""
    if world == 1:
        return None
    path, born = _id_path(tag)
    if rank == 0:
        try:
            os.unlink(path)                                     # (a crashed earlier job's: never ours)
        except OSError:
            pass
        if make_id is None:
            buf = ctypes.create_string_buffer(ID_BYTES)
            rc = lib().krrt_mgpu_unique_id(buf)
            if rc != KRRT_OK:
                raise KrrtError(rc, "krrt_mgpu_unique_id failed")
            raw = buf.raw
        else:
            raw = make_id()
        tmp = path + f".{os.getpid()}"
        with open(tmp, "wb") as f:
            f.write(raw)
        os.replace(tmp, path)
        return raw
    t0 = time.time()
    not_before = born - 2.0 if born is not None else t0 - 600.0  # (2 s: /proc start times are in clock ticks, mtimes are not)
    while True:
        try:
            if os.path.getmtime(path) >= not_before:
                with open(path, "rb") as f:
                    data = f.read()
                if len(data) == ID_BYTES:
                    return data
        except OSError:
            pass
        if time.time() - t0 > timeout:
            raise KrrtError(-1, f"no NCCL id at {path} after {timeout} s")
        time.sleep(0.01)


def release_id(rank, world, tag=None):
    """After the communicator exists on rank 0 (ncclCommInitRank is collective: every rank has read the id by then)."""
    if world > 1 and rank == 0:
        try:
            os.unlink(_id_path(tag)[0])
        except OSError:
            pass


class Sweep:
    """One libkrrt_mgpu handle."""

    def __init__(self, handle, cfg):
        self.h = handle
        self.cfg = cfg

    @classmethod
    def single_process(cls, n_gpus, devices=None, cfg=None, **kw):
        cfg = cfg if cfg is not None else capi.default_config(**kw)
        h = ctypes.c_void_p()
        dv = None
        if devices is not None:
            dv = (ctypes.c_int32 * n_gpus)(*devices)
        rc = lib().krrt_mgpu_create(ctypes.byref(cfg), int(n_gpus), dv, ctypes.byref(h))
        if rc != KRRT_OK:
            raise KrrtError(rc, lib().krrt_mgpu_last_error(None).decode())
        return cls(h, cfg)

    @classmethod
    def from_env(cls, cfg=None, **kw):
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        local = int(os.environ.get("LOCAL_RANK", "0"))
        cfg = cfg if cfg is not None else capi.default_config(**kw)
        cfg.device = local
        ident = exchange_id(rank, world)
        h = ctypes.c_void_p()
        rc = lib().krrt_mgpu_create_rank(ctypes.byref(cfg), rank, world, ident, ctypes.byref(h))
        release_id(rank, world)
        if rc != KRRT_OK:
            raise KrrtError(rc, lib().krrt_mgpu_last_error(None).decode())
        return cls(h, cfg)

    def _ck(self, rc):
        if rc != KRRT_OK:
            raise KrrtError(rc, lib().krrt_mgpu_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            lib().krrt_mgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        v = [ctypes.c_int32(0) for _ in range(5)]
        self._ck(lib().krrt_mgpu_info(self.h, *[ctypes.byref(x) for x in v]))
        return dict(zip(("world", "local", "first", "nccl", "nccl_version"), (x.value for x in v)))

    # solve_rrt() x Q in one call (…node.cpp:167-196): host arrays in, all Q records out, global order
    def sweep(self, starts, goals, obstacles, dynamic, n_iter, query_base=0, out=None):
        starts = capi._f64(starts, (-1, 4))
        goals = capi._f64(goals, (-1, 4))
        q = starts.shape[0]
        obstacles = capi._f64(obstacles)
        obstacles = obstacles.reshape(q, -1, 4) if obstacles.size else np.zeros((q, 0, 4))
        m = obstacles.shape[1]
        if out is None:
            out = np.zeros(q, dtype=RESULT_DTYPE)
        self._ck(lib().krrt_sweep_multi_gpu(self.h, capi._d(starts), capi._d(goals), capi._d(obstacles), q, m,
                                            int(dynamic), int(n_iter), int(query_base), out.ctypes.data))
        return out

    def load(self, starts, goals, obstacles):
        starts = capi._f64(starts, (-1, 4))
        goals = capi._f64(goals, (-1, 4))
        q = starts.shape[0]
        obstacles = capi._f64(obstacles)
        obstacles = obstacles.reshape(q, -1, 4) if obstacles.size else np.zeros((q, 0, 4))
        self._q = q
        self._ck(lib().krrt_sweep_load(self.h, capi._d(starts), capi._d(goals), capi._d(obstacles), q,
                                       obstacles.shape[1]))

    def run(self, dynamic, n_iter, query_base=0):
        """plans the loaded sweep and gathers the records; returns the device time in ms (this handle's GPUs)"""
        ms = ctypes.c_double(0.0)
        self._ck(lib().krrt_sweep_run(self.h, int(dynamic), int(n_iter), int(query_base), ctypes.byref(ms)))
        return ms.value

    def results(self, out=None):
        if out is None:
            out = np.zeros(self._q, dtype=RESULT_DTYPE)
        self._ck(lib().krrt_sweep_results(self.h, out.ctypes.data))
        return out

    def stats(self):
        s = KrrtStats()
        self._ck(lib().krrt_sweep_stats(self.h, ctypes.byref(s)))
        return {n: getattr(s, n) for n, _ in KrrtStats._fields_}

    def last_timing(self):
        a, b = ctypes.c_double(0.0), ctypes.c_double(0.0)
        self._ck(lib().krrt_sweep_last_timing(self.h, ctypes.byref(a), ctypes.byref(b)))
        return {"run_ms": a.value, "e2e_ms": b.value}

    def overlap(self):
        a, b = ctypes.c_int32(0), ctypes.c_int32(0)
        self._ck(lib().krrt_sweep_overlap(self.h, ctypes.byref(a), ctypes.byref(b)))
        return {"overlapped": bool(a.value), "chunks": b.value}

    @property
    def launch_count(self):
        return int(lib().krrt_mgpu_launch_count(self.h))

    def barrier(self):
        self._ck(lib().krrt_mgpu_barrier(self.h))

    def max_over_ranks(self, x):
        v = ctypes.c_double(float(x))
        self._ck(lib().krrt_mgpu_max(self.h, ctypes.byref(v)))
        return v.value
