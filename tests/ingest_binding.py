"""ctypes binding of the obstacle-ingest checker (oracle/ingest_oracle.c) and of oracle/_ref/libingest_ref.so — the
REFERENCE's own robotsubscriber.cpp compiled where it lies (oracle/ref_wrap/ingest_ref_wrap.cpp).  Test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_build", "libingest_oracle.so")
D = ctypes.POINTER(ctypes.c_double)
I = ctypes.POINTER(ctypes.c_int32)
F = ctypes.POINTER(ctypes.c_float)
_fn = None


def _lib():
    global _fn
    if _fn is None:
        src = os.path.join(ROOT, "oracle", "ingest_oracle.c")
        if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s", "all"])
        L = ctypes.CDLL(SO)
        _fn = L.io_ingest_vision
        _fn.argtypes = [F, I, ctypes.c_int, F, ctypes.c_int, I, ctypes.c_int, D, D, I, D, I]
        _fn.restype = None
    return _fn


REF_SO = os.path.join(ROOT, "oracle", "_ref", "libingest_ref.so")
_ref_fn = None


def have_ref():
    """built in the build container (where /root/reference is mounted); the .so travels to the GPU box"""
    if not os.path.exists(REF_SO) and os.path.isdir("/root/reference/src/robosoccer_motion_planning/src"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s", "ref"])
    return os.path.exists(REF_SO)


def _ref():
    global _ref_fn
    if _ref_fn is None:
        _ref_fn = ctypes.CDLL(REF_SO).ingest_ref_vision
        _ref_fn.argtypes = [F, I, ctypes.c_int, F, ctypes.c_int, I, ctypes.c_int, D, D, I, D, I]
        _ref_fn.restype = ctypes.c_int
    return _ref_fn


def reference_ingest(team, agent_id, rival, self_id):
    """the reference's RobotSubscriber fed with the same frames (one fresh subscriber per frame)"""
    return _ingest(_ref(), team, agent_id, rival, self_id)


def oracle_ingest(team, agent_id, rival, self_id):
    return _ingest(_lib(), team, agent_id, rival, self_id)


def _ingest(fn, team, agent_id, rival, self_id):
    team = np.ascontiguousarray(team, dtype=np.float32)
    rival = np.ascontiguousarray(rival, dtype=np.float32)
    agent_id = np.ascontiguousarray(agent_id, dtype=np.int32)
    self_id = np.ascontiguousarray(self_id, dtype=np.int32)
    f = self_id.shape[0]
    n_team = team.shape[1] if team.ndim == 3 else 0
    n_rival = rival.shape[1] if rival.ndim == 3 else 0
    out = {"self_state": np.zeros((f, 4)), "heading": np.zeros(f), "self_index": np.zeros(f, dtype=np.int32),
           "obstacles": np.zeros((f, n_team + n_rival, 4)), "n_obs": np.zeros(f, dtype=np.int32)}
    rc = fn(team.ctypes.data_as(F), agent_id.ctypes.data_as(I), n_team, rival.ctypes.data_as(F), n_rival,
           self_id.ctypes.data_as(I), f, out["self_state"].ctypes.data_as(D), out["heading"].ctypes.data_as(D),
           out["self_index"].ctypes.data_as(I), out["obstacles"].ctypes.data_as(D), out["n_obs"].ctypes.data_as(I))
    assert not rc, rc
    return out


def random_frames(rng, f, n_team=5, n_rival=5, p_missing=0.05, p_dup=0.02):
    """Synthetic OminiVisionInfo frames in nubot units: field 1800 x 1200 cm, speeds up to 300 cm/s."""
    team = np.empty((f, n_team, 5), dtype=np.float32)
    team[..., 0] = rng.uniform(-900, 900, (f, n_team))
    team[..., 1] = rng.uniform(-600, 600, (f, n_team))
    team[..., 2:4] = rng.uniform(-300, 300, (f, n_team, 2))
    team[..., 4] = rng.uniform(-np.pi, np.pi, (f, n_team))
    rival = np.empty((f, n_rival, 4), dtype=np.float32)
    rival[..., 0] = rng.uniform(-900, 900, (f, n_rival))
    rival[..., 1] = rng.uniform(-600, 600, (f, n_rival))
    rival[..., 2:4] = rng.uniform(-300, 300, (f, n_rival, 2))
    agent_id = np.tile(np.arange(1, n_team + 1, dtype=np.int32), (f, 1))
    agent_id = rng.permuted(agent_id, axis=1)
    self_id = rng.integers(1, max(2, n_team + 1), f).astype(np.int32)
    miss = rng.random(f) < p_missing
    self_id[miss] = 99                                  # the robot's own entry is absent from the message
    dup = rng.random(f) < p_dup
    if n_team >= 2:
        agent_id[dup, 0] = agent_id[dup, 1]             # a repeated AgentID: the later entry wins
    return team, agent_id, rival, self_id
