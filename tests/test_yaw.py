"""Yaw (heading) profiles, SURVEY.md §8(f) row 2: Trajectory1D::AngleController
(/root/reference/src/robosoccer_motion_planning/src/trajectory1d.cpp; test/unit_test.cpp:7-86).

PINNED parity chain:  reference library (oracle/_ref, built from the reference's own sources)
                      == committed fixtures generated from it (tests/golden/yaw_ref.npz)
                      == CPU restatement (oracle/yaw_oracle.c)        [CPU tests]
                      == CUDA kernel k_yaw_profiles through the C ABI  [GPU tests], all bit-for-bit.
"""
import os
import subprocess

import numpy as np
import pytest

import yaw_binding as yb

HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ("states", "times", "inputs", "final_time", "status", "n_seg")


@pytest.fixture(scope="module", autouse=True)
def _built():
    subprocess.check_call(["make", "-C", os.path.join(os.path.dirname(HERE), "oracle"), "-s", "all"])


def golden():
    return np.load(os.path.join(HERE, "golden", "yaw_ref.npz"))


def assert_same(a, b, what):
    for k in FIELDS:
        assert np.array_equal(a[k], b[k]), (what, k)


# ---- the reference's own known answers (test/unit_test.cpp) -------------------------------------------------------
UT_FROM = [np.pi + 0.1, 0.0, 1.32, -3.0, 1.7, 2.3, -2.3]
UT_TO = [np.pi, 1.32, 0.0, 1.7, -3.0, -1.4, 1.4]
UT_RES = [-0.1, 1.32, -1.32, -1.58, 1.58, 2.58, -2.58]


def test_shortest_path_known_answers():
    """TEST(AngleTrajectoryGenerator, angle_shortest_path), unit_test.cpp:7-32: |d - res| < 0.01."""
    for f, t, r in zip(UT_FROM, UT_TO, UT_RES):
        d, _, _ = yb.shortest_path("oracle", f, t)
        assert abs(d - r) < 0.01, (f, t, d)


def test_trajectory_known_answers():
    """TEST(AngleTrajectoryGenerator, trajectory), unit_test.cpp:34-86: 20 samples, every |w| <= pi and the last
    sample within 0.01 of the target (shortest angular distance)."""
    f = np.array([np.pi + 0.1, np.pi, 0.0, 1.32, -3.0, 1.7, 2.3, -2.3])
    t = np.array([np.pi, np.pi + 0.1, 1.32, 0.0, 1.7, -3.0, -1.4, 1.4])
    out = yb.oracle_profiles(f, np.zeros_like(f), t, 1.0, 1.0, 20)
    assert (out["status"] == 0).all()
    assert (np.abs(out["states"][:, :, 0]) <= np.pi).all()

    def angle_dist(a, b):
        d = a - b
        return d if abs(d) < np.pi else (1 if d < 0 else -1) * (2 * np.pi - (abs(a) + abs(b)))
    for k in range(len(f)):
        assert abs(angle_dist(out["states"][k, -1, 0], t[k])) <= 0.01, k


def test_restatement_equals_reference_fixtures():
    g = golden()
    for tag in "abc":
        vm, am, n = g[f"{tag}_params"]
        out = yb.oracle_profiles(g["w"], g["dw"], g["yaw"], vm, am, int(n))
        assert_same(out, {k: g[f"{tag}_{k}"] for k in FIELDS}, tag)


@pytest.mark.skipif(not yb.have_ref(), reason="oracle/_ref/libyaw_ref.so not built (needs /root/reference)")
@pytest.mark.parametrize("vm,am,n", [(1.0, 1.0, 25), (2.0, 0.7, 40), (0.5, 3.0, 7), (1.0, 1.0, 0), (3.0, 0.2, 1)])
def test_restatement_equals_reference_library(vm, am, n):
    rng = np.random.default_rng(int(vm * 100 + n))
    w, dw, yaw = yb.random_queries(rng, 50000)
    assert_same(yb.oracle_profiles(w, dw, yaw, vm, am, n), yb.ref_profiles(w, dw, yaw, vm, am, n), (vm, am, n))


def test_profile_properties():
    """Size-independent properties: |dw| never exceeds max(v_max, |dw0|), |effort| is 0 or a_max, time grids are
    uniform, the profile ends at rest on the target."""
    rng = np.random.default_rng(5)
    w, dw, yaw = yb.random_queries(rng, 20000)
    out = yb.oracle_profiles(w, dw, yaw, 1.0, 1.0, 32)
    ok = out["status"] == 0
    assert ok.all()
    assert (np.abs(out["states"][:, :, 1]) <= np.maximum(1.0, np.abs(dw))[:, None] + 1e-12).all()
    eff = np.abs(out["inputs"])
    assert (np.isclose(eff, 0.0) | np.isclose(eff, 1.0)).all()
    assert np.allclose(out["times"][:, -1], out["final_time"])
    moving = out["final_time"] > 0
    assert np.allclose(out["states"][moving, -1, 1], 0.0)
    d = np.abs(np.angle(np.exp(1j * (out["states"][moving, -1, 0] - yaw[moving]))))
    assert (d < 2e-3).all()   # the reference stops within 1e-5 of the target angle / 1e-3 of rest


# ---- CUDA kernel through the C ABI ----------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("vm,am,n", [(1.0, 1.0, 25), (2.0, 0.7, 40), (0.5, 3.0, 7), (1.0, 1.0, 0), (3.0, 0.2, 1)])
def test_gpu_equals_restatement(gpu_required, vm, am, n):
    from robosoccer_motion_planning_b200 import capi
    G = capi.Planner()
    rng = np.random.default_rng(int(vm * 100 + n) + 1)
    w, dw, yaw = yb.random_queries(rng, 100003)   # ragged: not a multiple of the warps per CTA
    assert_same(G.yaw_profiles(w, dw, yaw, vm, am, n), yb.oracle_profiles(w, dw, yaw, vm, am, n), (vm, am, n))
    G.close()


@pytest.mark.gpu
def test_gpu_equals_reference_fixtures(gpu_required):
    from robosoccer_motion_planning_b200 import capi
    G = capi.Planner()
    g = golden()
    for tag in "abc":
        vm, am, n = g[f"{tag}_params"]
        out = G.yaw_profiles(g["w"], g["dw"], g["yaw"], vm, am, int(n))
        assert_same(out, {k: g[f"{tag}_{k}"] for k in FIELDS}, tag)
    if yb.have_ref():   # the prebuilt reference library travels to the GPU box
        rng = np.random.default_rng(77)
        w, dw, yaw = yb.random_queries(rng, 30000)
        assert_same(G.yaw_profiles(w, dw, yaw, 1.0, 1.0, 25), yb.ref_profiles(w, dw, yaw, 1.0, 1.0, 25), "ref lib")
    G.close()


@pytest.mark.gpu
def test_gpu_edge_cases(gpu_required):
    from robosoccer_motion_planning_b200 import capi
    G = capi.Planner()
    out = G.yaw_profiles(np.zeros(0), np.zeros(0), np.zeros(0), 1.0, 1.0, 25)      # empty batch
    assert out["states"].shape == (0, 25, 2)
    w = np.array([0.5, np.nan, 0.0, 1.0])
    dw = np.array([0.0, 0.0, 0.0, 1.0])           # last: exactly v_max -> cruising case
    yaw = np.array([0.5, 1.0, np.nan, 3.0])       # first: already there; NaNs: the reference throws "case not match!"
    a = G.yaw_profiles(w, dw, yaw, 1.0, 1.0, 16)
    b = yb.oracle_profiles(w, dw, yaw, 1.0, 1.0, 16)
    assert_same(a, b, "edge")
    assert a["n_seg"][0] == 1 and a["final_time"][0] == 0.0
    assert a["status"][1] != 0 and a["status"][2] != 0
    with pytest.raises(capi.KrrtError):
        G.yaw_profiles(w, dw, yaw, -1.0, 1.0, 16)
    G.close()
