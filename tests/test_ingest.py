"""Obstacle ingest: OminiVisionInfo frames -> planner obstacle rows.
CPU: the restatement (oracle/ingest_oracle.c) against hand-computed known answers of robotsubscriber.cpp:29-77.
GPU: k_ingest_vision through the C ABI against the restatement, bit for bit, incl. feeding the batched planner."""
import numpy as np
import pytest

import ingest_binding as ib

KEYS = ("self_state", "heading", "self_index", "obstacles", "n_obs")


def test_known_answers():
    # one frame, 3 team robots (ids 2, 1, 3), robot 1 is "us"; 2 rivals
    team = np.array([[[150.0, -250.0, 30.0, -70.0, 0.5],
                      [-845.5, 12.25, 1.0, 2.0, -1.25],
                      [333.3, 0.1, -0.7, 299.9, 3.0]]], dtype=np.float32)
    ids = np.array([[2, 1, 3]], dtype=np.int32)
    rival = np.array([[[700.0, 100.0, -50.0, 25.0], [0.3, -0.7, 33.3, -299.9]]], dtype=np.float32)
    out = ib.oracle_ingest(team, ids, rival, np.array([1], dtype=np.int32))
    t64, r64 = team.astype(np.float64), rival.astype(np.float64)
    assert out["self_index"][0] == 1 and out["n_obs"][0] == 4
    # own state: float32 widened, then DIVIDED by 100. (robotsubscriber.cpp:51-54,62-63)
    assert np.array_equal(out["self_state"][0], t64[0, 1, :4] / 100.0)
    assert out["heading"][0] == np.float64(np.float32(-1.25))
    # friends in message order without self (robotsubscriber.cpp:70-77), then rivals flipped and MULTIPLIED by 1e-2
    # (robotsubscriber.cpp:36-39; getObstacles robotsubscriber.h:50-55)
    assert np.array_equal(out["obstacles"][0, 0], t64[0, 0, :4] / 100.0)
    assert np.array_equal(out["obstacles"][0, 1], t64[0, 2, :4] / 100.0)
    assert np.array_equal(out["obstacles"][0, 2], -r64[0, 0] * 1e-2)
    assert np.array_equal(out["obstacles"][0, 3], -r64[0, 1] * 1e-2)
    assert np.array_equal(out["obstacles"][0, 4], np.zeros(4))
    # exact decimal spot values: 150 cm -> 1.5 m, -250 -> -2.5, 700 -> -7.0 (flip)
    assert out["obstacles"][0, 0, 0] == 1.5 and out["obstacles"][0, 0, 1] == -2.5 and out["obstacles"][0, 2, 0] == -7.0
    # the two scalings are NOT the same rounding: some float32 value must tell x / 100. from x * 1e-2 apart
    x = np.arange(1, 2000, dtype=np.float32).astype(np.float64)
    assert np.any(x / 100.0 != x * 1e-2)


def test_absent_and_repeated_self():
    team = np.array([[[100, 200, 1, 2, 0.1], [300, 400, 3, 4, 0.2]]], dtype=np.float32)
    rival = np.zeros((1, 0, 4), dtype=np.float32)
    out = ib.oracle_ingest(team, np.array([[1, 2]], dtype=np.int32), rival, np.array([7], dtype=np.int32))
    assert out["self_index"][0] == -1 and out["n_obs"][0] == 2 and np.all(out["self_state"] == 0)
    assert np.array_equal(out["obstacles"][0], team[0, :, :4].astype(np.float64) / 100.0)
    out = ib.oracle_ingest(team, np.array([[5, 5]], dtype=np.int32), rival, np.array([5], dtype=np.int32))
    assert out["self_index"][0] == 1 and out["n_obs"][0] == 0          # the later entry overwrites the state
    assert np.array_equal(out["self_state"][0], np.array([3.0, 4.0, 0.03, 0.04]))


def test_vectorised_numpy_agrees():
    rng = np.random.default_rng(3)
    team, ids, rival, me = ib.random_frames(rng, 500, 5, 5, p_missing=0.0, p_dup=0.0)
    out = ib.oracle_ingest(team, ids, rival, me)
    assert np.all(out["n_obs"] == 9)
    for q in range(500):
        keep = ids[q] != me[q]
        exp = np.concatenate([team[q, keep, :4].astype(np.float64) / 100.0, -rival[q].astype(np.float64) * 1e-2])
        assert np.array_equal(out["obstacles"][q, :9], exp)


@pytest.mark.gpu
@pytest.mark.parametrize("f,n_team,n_rival", [(1, 5, 5), (1000, 5, 5), (257, 1, 0), (64, 0, 3), (4096, 11, 11), (3, 2, 60)])
def test_gpu_ingest_matches_oracle(gpu_required, f, n_team, n_rival):
    from robosoccer_motion_planning_b200 import capi
    rng = np.random.default_rng(f + n_team)
    team, ids, rival, me = ib.random_frames(rng, f, n_team, n_rival)
    G = capi.Planner()
    l0 = G.launch_count
    got = G.ingest_vision(team, ids, rival, me)
    assert G.launch_count == l0 + 1
    ref = ib.oracle_ingest(team, ids, rival, me)
    for k in KEYS:
        assert np.array_equal(got[k], ref[k]), k
    G.close()


@pytest.mark.gpu
def test_gpu_ingest_feeds_batched_planner(gpu_required):
    """Recorded-style frames drive the batched planner instead of synthetic scenes: the rows the ingest kernel wrote are
    the rows krrt_plan_batch takes (5 v 5 -> 9 obstacles, the reference's n_obstacles, robotsubscriber.h:74)."""
    import oracle_binding as ob
    from robosoccer_motion_planning_b200 import capi
    rng = np.random.default_rng(11)
    q = 64
    team, ids, rival, me = ib.random_frames(rng, q, 5, 5, p_missing=0.0, p_dup=0.0)
    G = capi.Planner(p_direct=0.1)
    got = G.ingest_vision(team, ids, rival, me)
    assert np.all(got["n_obs"] == 9)
    starts = np.ascontiguousarray(got["self_state"])
    goals = np.zeros((q, 4))
    goals[:, 0] = rng.uniform(-8, 8, q)
    goals[:, 1] = rng.uniform(-5, 5, q)
    obstacles = np.ascontiguousarray(got["obstacles"][:, :9])
    out = G.plan_batch(starts, goals, obstacles, 1, 60)
    ref_in = ib.oracle_ingest(team, ids, rival, me)
    ref = ob.plan_batch(ob.default_config(p_direct=0.1), np.ascontiguousarray(ref_in["self_state"]), goals,
                        np.ascontiguousarray(ref_in["obstacles"][:, :9]), 1, 60, nthreads=4)
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
    G.close()


# ---- pinned against the REFERENCE's own code: robotsubscriber.cpp compiled where it lies (oracle/_ref/libingest_ref.so) ----
REF_SHAPES = [(1, 5, 5), (1000, 5, 5), (257, 1, 0), (64, 0, 3), (512, 11, 11), (3, 2, 60)]


def _need_ref():
    if not ib.have_ref():
        pytest.skip("oracle/_ref/libingest_ref.so is built where /root/reference is mounted (make -C oracle ref)")


@pytest.mark.parametrize("f,n_team,n_rival", REF_SHAPES)
def test_restatement_equals_reference_subscriber(f, n_team, n_rival):
    """oracle/ingest_oracle.c against RobotSubscriber::callback / rival_callback / getObstacles themselves, bit for bit:
    frames with the own entry missing, with a repeated AgentID, with no friends, with no rivals"""
    _need_ref()
    rng = np.random.default_rng(100 + f + n_team)
    team, ids, rival, me = ib.random_frames(rng, f, n_team, n_rival, p_missing=0.1, p_dup=0.1)
    ref = ib.reference_ingest(team, ids, rival, me)
    got = ib.oracle_ingest(team, ids, rival, me)
    for k in KEYS:
        assert np.array_equal(got[k], ref[k]), k
    if f >= 257 and n_team >= 5:
        assert (ref["self_index"] < 0).any() and (ref["n_obs"] < n_team + n_rival - 1).any()   # the odd frames are in there


def test_reference_subscriber_known_answers():
    """the hand-computed answers of test_known_answers, asked of the reference's code"""
    _need_ref()
    team = np.array([[[150.0, -250.0, 30.0, -70.0, 0.5], [-845.5, 12.25, 1.0, 2.0, -1.25], [333.3, 0.1, -0.7, 299.9, 3.0]]],
                    dtype=np.float32)
    rival = np.array([[[700.0, 100.0, -50.0, 25.0], [0.3, -0.7, 33.3, -299.9]]], dtype=np.float32)
    out = ib.reference_ingest(team, np.array([[2, 1, 3]], dtype=np.int32), rival, np.array([1], dtype=np.int32))
    t64, r64 = team.astype(np.float64), rival.astype(np.float64)
    assert out["self_index"][0] == 1 and out["n_obs"][0] == 4
    assert np.array_equal(out["self_state"][0], t64[0, 1, :4] / 100.0) and out["heading"][0] == np.float64(np.float32(-1.25))
    assert np.array_equal(out["obstacles"][0, :2], t64[0, [0, 2], :4] / 100.0)
    assert np.array_equal(out["obstacles"][0, 2:4], -r64[0] * 1e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("f,n_team,n_rival", REF_SHAPES)
def test_gpu_ingest_equals_reference_subscriber(gpu_required, f, n_team, n_rival):
    """k_ingest_vision through the C ABI against the reference's subscriber (the library built in the build container
    travels to the GPU box; nothing under /root/reference is read here)"""
    from robosoccer_motion_planning_b200 import capi
    assert ib.have_ref(), "oracle/_ref/libingest_ref.so did not travel: run __graft_entry__.build() where /root/reference is mounted"
    rng = np.random.default_rng(100 + f + n_team)
    team, ids, rival, me = ib.random_frames(rng, f, n_team, n_rival, p_missing=0.1, p_dup=0.1)
    G = capi.Planner()
    got = G.ingest_vision(team, ids, rival, me)
    ref = ib.reference_ingest(team, ids, rival, me)
    for k in KEYS:
        assert np.array_equal(got[k], ref[k]), k
    G.close()
