"""GPU parity tests proper: libkrrt.so (through its C ABI) against the CPU oracle on the same seeded inputs.

Bar: bit-exact — states, parents, costs, arrival times, indices, counters (fp64 mode computes with the same
IEEE operations in the same order as the oracle, so `==` is the tolerance).  north_star would allow 1e-6
relative on costs/times and ties within 1e-9; we do not need the allowance.
"""
import ctypes

import numpy as np
import pytest

import oracle_binding as ob
from robosoccer_motion_planning_b200 import capi, scenes

pytestmark = pytest.mark.gpu


def rand_states(rng, n):
    return rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, (n, 4))


def make_pair(gpu_required, dynamic, p_direct=0.0, capacity=1 << 15, batch=0, near_cap=0, seed=0, query_id=0, **kw):
    common = dict(capacity=capacity, p_direct=p_direct, seed=seed, query_id=query_id, **kw)
    G = capi.Planner(batch=batch, near_cap=near_cap, **common)
    O = ob.OraclePlanner(**common)
    obs = scenes.dynamic_obstacles() if dynamic else scenes.static_obstacles()
    for P in (G, O):
        P.set_obstacles(obs, dynamic)
        P.set_start(scenes.START)
    return G, O


def assert_same_tree(G, O):
    gs, gp, gc, gt = G.tree()
    os_, op, oc, ot = O.tree()
    assert G.size == O.size
    assert np.array_equal(gp, op), f"parents differ at {np.nonzero(gp != op)[0][:10]}"
    assert np.array_equal(gs, os_)
    assert np.array_equal(gc, oc), f"cost-to-come differs at {np.nonzero(gc != oc)[0][:10]}"
    assert np.array_equal(gt, ot)
    assert G.iteration == O.iteration
    assert G.stats() == O.stats()


def test_steer_cost_bit_exact(gpu_required):
    rng = np.random.default_rng(0)
    n = 200_000
    a, b = rand_states(rng, n), rand_states(rng, n)
    b[::7, :2] = a[::7, :2] + 1e-3 * rng.uniform(-1, 1, (len(a[::7]), 2))   # near-coincident
    a[::11, 2:] = 0.0; b[::11, 2:] = 0.0                                     # rest to rest
    b[::13, :2] = a[::13, :2]                                                # same position
    a[::17, 2:] *= 4; b[::17, 2:] *= 4                                       # multi-root zone (fast, close)
    b[::17, :2] = a[::17, :2] + 0.3 * rng.uniform(-1, 1, (len(a[::17]), 2))
    G = capi.Planner(capacity=16)
    t, c = G.steer_cost(a, b)
    to, co = ob.steer(a, b, 0.5)
    assert np.array_equal(t, to) and np.array_equal(c, co)
    # empty and single-element inputs
    t0, c0 = G.steer_cost(np.zeros((0, 4)), np.zeros((0, 4)))
    assert t0.shape == (0,) and c0.shape == (0,)
    t1, c1 = G.steer_cost(a[:1], b[:1])
    assert t1[0] == to[0] and c1[0] == co[0]


def test_steer_cost_other_weight(gpu_required):
    rng = np.random.default_rng(5)
    a, b = rand_states(rng, 5000), rand_states(rng, 5000)
    for r in (0.05, 1.0, 7.5):
        G = capi.Planner(capacity=16, r=r)
        t, c = G.steer_cost(a, b)
        to, co = ob.steer(a, b, r)
        assert np.array_equal(t, to) and np.array_equal(c, co)


@pytest.mark.parametrize("dynamic", [0, 1])
def test_edge_collide_bit_exact(gpu_required, dynamic):
    rng = np.random.default_rng(1)
    n = 20_000
    a, b = rand_states(rng, n), rand_states(rng, n)
    tau, _ = ob.steer(a, b, 0.5)
    t0 = rng.uniform(0, 10, n)
    obs = scenes.dynamic_obstacles(3)
    G = capi.Planner(capacity=16)
    G.set_obstacles(obs, dynamic)
    hit = G.edge_collide(a, b, tau, t0)
    ref = np.array([ob.edge_collide(a[i], b[i], tau[i], t0[i], obs, dynamic) for i in range(n)], dtype=np.uint8)
    assert np.array_equal(hit, ref)
    assert 0 < hit.sum() < n
    # no obstacles -> never hits
    G.set_obstacles(np.zeros((0, 4)), dynamic)
    assert G.edge_collide(a[:100], b[:100], tau[:100], t0[:100]).sum() == 0


def test_solve_rows_bit_exact(gpu_required):
    rng = np.random.default_rng(2)
    G = capi.Planner(capacity=16)
    for k in range(40):
        a, b = rand_states(rng, 1)[0], rand_states(rng, 1)[0]
        dt = [0.1, 0.05, 0.37, 5.0][k % 4]
        rows = G.solve(a, b, dt=dt, t0=1.25)
        ref = ob.solve(a, b, 0.5, dt, 1.25)
        assert rows.shape == ref.shape
        assert np.array_equal(rows, ref)
    # identical states with a velocity: a genuine loop of duration sqrt(12 r |v|^2)
    rows = G.solve(a, a, dt=0.1)
    ref = ob.solve(a, a, 0.5, 0.1, 0.0)
    assert rows.shape == ref.shape and rows.shape[0] > 1 and np.array_equal(rows, ref)
    # degenerate: identical states at rest -> tau = 0, a single row
    z = np.array([a[0], a[1], 0.0, 0.0])
    rows = G.solve(z, z, dt=0.1)
    ref = ob.solve(z, z, 0.5, 0.1, 0.0)
    assert rows.shape == ref.shape == (1, 5) and np.array_equal(rows, ref)


@pytest.mark.parametrize("batch", [1, 5, 64, 256])
def test_grow_static_c1(gpu_required, batch):
    G, O = make_pair(gpu_required, dynamic=0, batch=batch)
    for P in (G, O):
        P.grow(scenes.GOAL, 3000)
    assert_same_tree(G, O)
    assert O.check_invariants() == 0
    assert O.stats()["fail_reject"] > 0 and O.stats()["rewired"] > 0


@pytest.mark.parametrize("batch", [1, 64])
def test_grow_dynamic_c2(gpu_required, batch):
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, batch=batch)
    rg = G.grow(scenes.GOAL, 5000)
    ro = O.grow(scenes.GOAL, 5000)
    assert rg == ro and ro[0]
    assert_same_tree(G, O)
    assert O.stats()["fail_dup"] > 0
    a = G.result(*rg)
    b = O.result(*ro)
    for f in a.dtype.names:
        assert np.array_equal(a[f], b[f]), f
    assert np.array_equal(G.last_sample(), O.last_sample())


def test_grow_full_c2_20k(gpu_required):
    """BASELINE config 2 at full size: 9 moving opponents, p_direct 0.1, 20k iterations."""
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, capacity=20002)
    rg = G.grow(scenes.GOAL, 20000)
    ro = O.grow(scenes.GOAL, 20000)
    assert rg == ro
    assert_same_tree(G, O)
    ps, pt = G.path(rg[1])
    qs, qt = O.path(ro[1])
    assert np.array_equal(ps, qs) and np.array_equal(pt, qt)


def test_grow_incremental_calls_match_single_call(gpu_required):
    """The reference calls grow(xg) once per iteration (…node.cpp:176-177)."""
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1)
    for _ in range(150):
        G.grow(scenes.GOAL, 1)
    G.grow(scenes.GOAL, 77)
    O.grow(scenes.GOAL, 227)
    assert_same_tree(G, O)


def test_grow_explicit_samples_and_nan_rows(gpu_required):
    """sampler.set_next_sample(x) (…node.cpp:183-184): forced samples, NaN rows fall back to Philox."""
    rng = np.random.default_rng(3)
    n = 500
    samples = rand_states(rng, n)
    samples[::3, 0] = np.nan
    samples[250] = scenes.GOAL
    samples[498] = scenes.GOAL          # second visit of the goal state: duplicate -> failed iteration
    samples[11] = samples[10]           # duplicate inside one batch (rows 10, 11 are not NaN rows)
    G, O = make_pair(gpu_required, dynamic=1, batch=64)
    rg = G.grow(scenes.GOAL, n, samples)
    ro = O.grow(scenes.GOAL, n, samples)
    assert rg == ro and ro[0]
    assert_same_tree(G, O)
    assert O.stats()["fail_dup"] >= 2


def test_grow_no_goal_no_obstacles(gpu_required):
    G = capi.Planner(capacity=4096)
    O = ob.OraclePlanner(capacity=4096)
    for P in (G, O):
        P.set_obstacles(np.zeros((0, 4)), 0)
        P.set_start(scenes.START)
    rg = G.grow(None, 2000)
    ro = O.grow(None, 2000)
    assert rg == ro == (False, -1)
    assert_same_tree(G, O)


def test_capacity_full(gpu_required):
    G, O = make_pair(gpu_required, dynamic=1, capacity=300, batch=64)
    for P in (G, O):
        P.grow(scenes.GOAL, 500)
    done, full = G.grow_done()               # the call returns KRRT_OK; how it ended is on record
    assert full and done < 500 and done == O.iteration
    it0 = O.iteration
    for P in (G, O):
        P.grow(scenes.GOAL, 3)
    assert G.grow_done() == (O.iteration - it0, True)   # (the iteration that finds the tree full is consumed, on both sides)
    assert G.size == O.size == 300
    assert_same_tree(G, O)


def test_restart_reuses_context(gpu_required):
    """set_start + setIteration(0) every planning cycle (…node.cpp:118-137)."""
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1)
    for P in (G, O):
        P.grow(scenes.GOAL, 300)
        P.set_obstacles(scenes.dynamic_obstacles(9), 1)
        P.set_start([0.0, 1.0, 0.5, -0.5])
        P.grow([3.0, -2.0, 0.0, 0.0], 400)
    assert_same_tree(G, O)


def test_near_overflow_is_loud(gpu_required):
    G = capi.Planner(capacity=4096, near_cap=32, gamma=50.0, rho_max=50.0)
    G.set_obstacles(np.zeros((0, 4)), 0)
    G.set_start(scenes.START)
    with pytest.raises(capi.KrrtError) as e:
        G.grow(None, 500)
    assert e.value.code == capi.KRRT_ERR_NEAR_OVERFLOW


def test_near_query(gpu_required):
    G, O = make_pair(gpu_required, dynamic=0)
    for P in (G, O):
        P.grow(None, 4000)
    rng = np.random.default_rng(4)
    for k in range(20):
        x = rand_states(rng, 1)[0]
        r2 = [0.0, 0.3, 1.0, 9.0, 400.0][k % 5]
        gi, gn, gd = G.near(x, r2, cap=8192)
        oi, on, od = O.near(x, r2, cap=8192)
        assert np.array_equal(gi, oi) and gn == on and gd == od
    # a tree node itself: distance 0, radius 0 -> exactly that node
    s = O.tree()[0][123]
    gi, gn, gd = G.near(s, 0.0)
    assert list(gi) == [123] and gn == 123 and gd == 0.0


def test_trajectory_of_path(gpu_required):
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1)
    rg = G.grow(scenes.GOAL, 3000)
    ro = O.grow(scenes.GOAL, 3000)
    assert rg == ro and rg[0]
    st, tm = O.path(ro[1])
    ref = np.concatenate([ob.solve(st[k], st[k + 1], 0.5, 0.1, tm[k]) for k in range(len(st) - 1)])
    rows = G.trajectory(rg[1], 0.1)
    assert rows.shape == ref.shape and np.array_equal(rows, ref)
    # arrival time of the last row == arrival time of the goal node; last row == goal state (1e-9)
    assert abs(rows[-1, 4] - tm[-1]) <= 1e-9 * max(1.0, tm[-1])
    assert np.allclose(rows[-1, :4], st[-1], atol=1e-9)


@pytest.mark.parametrize("dynamic,n_iter,q", [(1, 200, 300), (0, 75, 200), (1, 1, 40), (1, 0, 8),
                                              (1, 30, 800), (1, 50, 500), (0, 25, 400), (1, 62, 300)])
def test_plan_batch(gpu_required, dynamic, n_iter, q):
    """(trees of fewer than 64 slots once lost the grandchildren of rewired nodes in the propagation queue: the small
    sizes of the live system — target_size 50 / 75, config.yaml:3, debug_core.launch:2 — are cases of their own)"""
    starts, goals, obs = scenes.batch_queries(q, 9, seed=11)
    cfg_kw = dict(p_direct=0.1, seed=5)
    G = capi.Planner(capacity=1 << 12, **cfg_kw)
    out = G.plan_batch(starts, goals, obs, dynamic, n_iter, query_base=1000)
    ref = ob.plan_batch(ob.default_config(capacity=1 << 12, **cfg_kw), starts, goals, obs, dynamic, n_iter,
                        query_base=1000, nthreads=8)
    for f in out.dtype.names:
        assert np.array_equal(out[f], ref[f]), f
    if n_iter >= 75:
        assert out["solved"].sum() > 0


def test_plan_batch_no_obstacles_and_stats(gpu_required):
    q, n_iter = 64, 120
    starts, goals, _ = scenes.batch_queries(q, 1, seed=2)
    obs = np.zeros((q, 0, 4))
    G = capi.Planner(p_direct=0.25)
    out = G.plan_batch(starts, goals, obs, 1, n_iter)
    cfg = ob.default_config(p_direct=0.25)
    ref = ob.plan_batch(cfg, starts, goals, obs, 1, n_iter, nthreads=4)
    for f in out.dtype.names:
        assert np.array_equal(out[f], ref[f]), f
    # counters: sum over per-query oracle planners
    tot = None
    for i in range(q):
        P = ob.OraclePlanner(p_direct=0.25, capacity=n_iter + 2, query_id=i)
        P.set_obstacles(np.zeros((0, 4)), 1)
        P.set_start(starts[i])
        P.grow(goals[i], n_iter)
        s = P.stats()
        tot = s if tot is None else {k: tot[k] + s[k] for k in s}
    assert G.batch_stats() == tot


def _oracle_batch_stats(starts, goals, obs, dynamic, n_iter, **cfg_kw):
    tot = None
    for i in range(len(starts)):
        P = ob.OraclePlanner(capacity=n_iter + 2, query_id=i, **cfg_kw)
        P.set_obstacles(obs[i], dynamic)
        P.set_start(starts[i])
        P.grow(goals[i], n_iter)
        s = P.stats()
        tot = s if tot is None else {k: tot[k] + s[k] for k in s}
    return tot


def test_plan_batch_near_overflow_takes_the_redo_pass(gpu_required):
    """A radius that makes every node a neighbour: near sets pass 64 entries (what k_plan_batch2 keeps in registers), so the
    queries are handed to k_plan_batch through the redo list — results and counters must not show it."""
    q, n_iter = 48, 130
    starts, goals, obs = scenes.batch_queries(q, 9, seed=4)
    kw = dict(p_direct=0.1, gamma=60.0, rho_max=40.0)
    G = capi.Planner(**kw)
    out = G.plan_batch(starts, goals, obs, 1, n_iter)
    ref = ob.plan_batch(ob.default_config(**kw), starts, goals, obs, 1, n_iter, nthreads=8)
    for f in out.dtype.names:
        assert np.array_equal(out[f], ref[f]), f
    tot = _oracle_batch_stats(starts, goals, obs, 1, n_iter, **kw)
    assert tot["sum_near"] > 64 * q            # the near sets did outgrow the registers
    assert G.batch_stats() == tot


@pytest.mark.parametrize("dynamic,S,m", [(1, 10, 9), (0, 10, 9), (1, 7, 3), (1, 40, 2), (1, 1, 5), (1, 70, 1), (1, 16, 64)])
def test_plan_batch_edge_test_lane_maps(gpu_required, dynamic, S, m):
    """k_plan_batch2 tests an edge with lane = (sample point, obstacle group): every split of the 32 lanes (S x m below,
    at, and above 32; S above the h table) must give the sequential planner's trees and counters."""
    q, n_iter = 32, 90
    starts, goals, obs = scenes.batch_queries(q, m, seed=9)
    kw = dict(p_direct=0.1, S=S)
    G = capi.Planner(**kw)
    out = G.plan_batch(starts, goals, obs, dynamic, n_iter)
    ref = ob.plan_batch(ob.default_config(**kw), starts, goals, obs, dynamic, n_iter, nthreads=8)
    for f in out.dtype.names:
        assert np.array_equal(out[f], ref[f]), f
    assert G.batch_stats() == _oracle_batch_stats(starts, goals, obs, dynamic, n_iter, **kw)


def test_plan_batch_both_kernels_agree(gpu_required, monkeypatch):
    """KRRT_BATCH_V1=1 plans the whole batch with k_plan_batch (lists in shared memory); the default is k_plan_batch2."""
    q, n_iter = 500, 200
    starts, goals, obs = scenes.batch_queries(q, 9, seed=21)
    G2 = capi.Planner(p_direct=0.1)
    out2 = G2.plan_batch(starts, goals, obs, 1, n_iter, query_base=77)
    st2 = G2.batch_stats()
    monkeypatch.setenv("KRRT_BATCH_V1", "1")
    G1 = capi.Planner(p_direct=0.1)
    monkeypatch.delenv("KRRT_BATCH_V1")
    out1 = G1.plan_batch(starts, goals, obs, 1, n_iter, query_base=77)
    assert np.array_equal(out1.view(np.uint8), out2.view(np.uint8))
    assert G1.batch_stats() == st2
    assert G1.launch_count == 1 and G2.launch_count == 2     # k_plan_batch | k_plan_batch2 + the (empty) redo pass


def test_large_tree_100k_properties_and_parity(gpu_required):
    """BASELINE config 3 (single large tree, 100k iterations): size-independent invariants on the GPU tree,
    then full bit parity (the oracle's kd-tree finishes this in a few seconds)."""
    n_iter = 100_000
    G, O = make_pair(gpu_required, dynamic=0, capacity=n_iter + 2)
    G.grow(None, n_iter)
    st, par, ctc, tim = G.tree()
    n = len(par)
    assert par[0] == -1 and ctc[0] == 0.0 and tim[0] == 0.0
    assert np.all((par[1:] >= 0) & (par[1:] < n))
    # cost / time telescoping with edge costs recomputed by the GPU steering operator
    et, ec = G.steer_cost(st[par[1:]], st[1:])
    assert np.array_equal(ctc[1:], ctc[par[1:]] + ec)
    assert np.array_equal(tim[1:], tim[par[1:]] + et)
    # acyclic: repeated parent jumps reach the root
    cur = np.arange(n)
    for _ in range(4096):
        cur = np.where(cur > 0, par[np.maximum(cur, 0)], cur)
        cur = np.maximum(cur, 0)
        if not cur.any():
            break
    assert not cur.any()
    O.grow(None, n_iter)
    assert_same_tree(G, O)


@pytest.mark.slow
def test_large_tree_1m_iterations_parity(gpu_required):
    """BASELINE config 3 at its upper end: 1 M iterations (737 k nodes; grid index rebuilt ~60 times, windows of 1024 samples,
    2 700 samples on the in-order path because they rewire nodes above > 1024 descendants).  Full bit parity: states, parents, costs,
    times, all ten counters.  The oracle's kd-tree needs about half a minute for this plan."""
    n_iter = 1_000_000
    G, O = make_pair(gpu_required, dynamic=0, capacity=n_iter + 2)
    G.grow(None, n_iter)
    O.grow(None, n_iter)
    assert G.size > 700_000
    assert_same_tree(G, O)


def test_edge_samples_and_tree_edges(gpu_required):
    """Connector slot: S samples per optimal edge, and tree.trajectories (robosoccer_visual.hpp:18)."""
    rng = np.random.default_rng(9)
    n, S = 3000, 10
    a, b = rand_states(rng, n), rand_states(rng, n)
    t0 = rng.uniform(0, 5, n)
    G, O = make_pair(gpu_required, 1, p_direct=0.1, capacity=2048)
    rows = G.edge_samples(a, b, t0, S)
    tau, _ = ob.steer(a, b, 0.5)
    for k in range(S):
        h = k / (S - 1)
        ref = ob.edge_state(a, b, tau, np.full(n, h))
        assert np.array_equal(rows[:, k, :4], ref)
        assert np.array_equal(rows[:, k, 4], t0 + tau * h)
    G.grow(scenes.GOAL, 1500)
    O.grow(scenes.GOAL, 1500)
    st, par, _, tim = O.tree()
    E = G.tree_edges(S)
    assert E.shape == (O.size, S, 5)
    assert np.array_equal(E[0, :, :4], np.tile(st[0], (S, 1))) and not E[0, :, 4].any()
    i = np.arange(1, O.size)
    tau, _ = ob.steer(st[par[i]], st[i], 0.5)
    for k in range(S):
        h = k / (S - 1)
        assert np.array_equal(E[1:, k, :4], ob.edge_state(st[par[i]], st[i], tau, np.full(len(i), h)))
        assert np.array_equal(E[1:, k, 4], tim[par[i]] + tau * h)
    assert np.array_equal(E[1:, S - 1, 4], tim[par[i]] + tau)


def test_set_params_equals_fresh_context(gpu_required):
    """solver.set_weight(r) / env.collision_radius / sampler settings changed on a live context (…node.cpp:84-92)."""
    kw = dict(r=1.25, R=0.9, gamma=5.0, rho_max=3.0, p_direct=0.2, seed=7, query_id=3, goal_tol=0.25)
    G = capi.Planner(capacity=4096)
    G.set_params(**kw)
    O = ob.OraclePlanner(capacity=4096, **kw)
    obs = scenes.dynamic_obstacles(2)
    for P in (G, O):
        P.set_obstacles(obs, 1)
        P.set_start(scenes.START)
    assert G.grow(scenes.GOAL, 1500) == O.grow(scenes.GOAL, 1500)
    assert_same_tree(G, O)
    with pytest.raises(capi.KrrtError):
        G.set_params(capacity=8192)


@pytest.mark.parametrize("dynamic,batch,min_n", [(0, 0, 300), (1, 0, 300), (1, 37, 1000), (0, 512, 2100)])
def test_grid_index_small_threshold(gpu_required, monkeypatch, dynamic, batch, min_n):
    """The spatial index of the neighbour scan (uniform grid over the position plane, rebuilt as the tree grows) switched
    on for small trees, so a 6000-iteration plan sees many rebuilds: same tree, same counters, fewer pairs visited."""
    monkeypatch.setenv("KRRT_GRID_MIN_N", str(min_n))
    G, O = make_pair(gpu_required, dynamic=dynamic, p_direct=0.1 if dynamic else 0.0, batch=batch)
    monkeypatch.delenv("KRRT_GRID_MIN_N")
    for k in range(3):                       # three calls: the index survives from one krrt_grow call to the next
        assert G.grow(scenes.GOAL, 2000) == O.grow(scenes.GOAL, 2000)
    assert_same_tree(G, O)
    ss = G.scan_stats()
    assert ss["index_rebuilds"] >= (2 if min_n <= 1000 else 1) and 0 < ss["indexed_nodes"] <= G.size
    assert ss["visited_pairs"] < 0.7 * G.last_timing()["scan_node_samples"]


def test_grid_index_off_is_the_same_tree(gpu_required, monkeypatch):
    monkeypatch.setenv("KRRT_GRID_INDEX", "0")
    G, O = make_pair(gpu_required, dynamic=0, capacity=30002)
    monkeypatch.delenv("KRRT_GRID_INDEX")
    G.grow(None, 30000)
    O.grow(None, 30000)
    assert_same_tree(G, O)
    ss = G.scan_stats()
    assert ss["index_rebuilds"] == 0 and ss["visited_pairs"] == G.last_timing()["scan_node_samples"]


def test_grid_index_survives_a_larger_radius(gpu_required, monkeypatch):
    """krrt_set_params raises the near radius on a live tree: the cells of the index are too small for it, so the next
    window rebuilds it (a stale index would silently drop neighbours)."""
    monkeypatch.setenv("KRRT_GRID_MIN_N", "500")
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, gamma=4.0, rho_max=2.0)
    monkeypatch.delenv("KRRT_GRID_MIN_N")
    assert G.grow(scenes.GOAL, 3000) == O.grow(scenes.GOAL, 3000)
    r0 = G.scan_stats()["index_rebuilds"]
    assert r0 >= 1
    G.set_params(gamma=7.0, rho_max=4.0)
    O.set_params(gamma=7.0, rho_max=4.0)
    assert G.grow(scenes.GOAL, 200) == O.grow(scenes.GOAL, 200)
    assert G.scan_stats()["index_rebuilds"] > r0
    assert G.grow(scenes.GOAL, 1500) == O.grow(scenes.GOAL, 1500)
    assert_same_tree(G, O)


@pytest.mark.parametrize("no_small", [False, True])
def test_single_expansion_calls_cross_the_small_tree_launch(gpu_required, monkeypatch, no_small):
    """The node's call pattern: one krrt_grow per iteration (…node.cpp:176-185).  While the tree stays below the size where
    sub-rounds start, a call is ONE CTA in a plain launch; past it, the cooperative grid.  Same tree either way, and the
    getters between two calls are served from one read-back."""
    if no_small:
        monkeypatch.setenv("KRRT_NO_SMALL", "1")
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, capacity=2048)
    for k in range(330):
        step = 1 if k % 7 else 3
        assert G.grow(scenes.GOAL, step) == O.grow(scenes.GOAL, step)
        if k % 50 == 0:
            assert G.size == O.size and np.array_equal(G.last_sample(), O.last_sample())
            assert G.iteration == O.iteration
    assert G.size > 256
    assert_same_tree(G, O)


@pytest.mark.parametrize("pend,raw", [("0", "0"), ("1", "0"), ("0", "1")])
def test_prefix_relaxations_can_be_switched_off(gpu_required, monkeypatch, pend, raw):
    """KRRT_PEND / KRRT_RAW_FORWARD (read at every krrt_grow call) switch off the two ways a sub-round's validated prefix is
    extended (same-sub-round neighbours settled at validation; rewired near nodes re-checked with their new cost).  The
    tree must not depend on them — they only change how many expansions a sub-round commits."""
    monkeypatch.setenv("KRRT_PEND", pend)
    monkeypatch.setenv("KRRT_RAW_FORWARD", raw)
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1)
    assert G.grow(scenes.GOAL, 8000) == O.grow(scenes.GOAL, 8000)
    assert_same_tree(G, O)
    off = G.wave_stats()["rounds_parallel"]
    monkeypatch.setenv("KRRT_PEND", "1")
    monkeypatch.setenv("KRRT_RAW_FORWARD", "1")
    G2, _ = make_pair(gpu_required, dynamic=1, p_direct=0.1)
    G2.grow(scenes.GOAL, 8000)
    assert G2.wave_stats()["rounds_parallel"] < off
    for a, b in zip(G.tree(), G2.tree()):
        assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct,n_iter", [(1, 0.1, 20000), (0, 0.0, 60000)])
def test_candidates_below_rewired_nodes(gpu_required, monkeypatch, dynamic, p_direct, n_iter):
    """A sample that rewires a node AND a candidate child of that node: settled inside decide (KRRT_INTRA_FIX, the child is
    re-tested at the cost the first rewire leaves it), else committed in order right behind the prefix (KRRT_TAIL_SEQ), else
    in a sub-round of its own.  Tree AND counters (collision tests, propagated nodes) equal the oracle's in all three modes;
    the switches only change the number of sub-rounds."""
    goal = scenes.GOAL if dynamic else None
    rounds = []
    for fix, tail in (("1", "1"), ("0", "1"), ("0", "0")):
        monkeypatch.setenv("KRRT_INTRA_FIX", fix)
        monkeypatch.setenv("KRRT_TAIL_SEQ", tail)
        G, O = make_pair(gpu_required, dynamic=dynamic, p_direct=p_direct, capacity=n_iter + 2)
        assert G.grow(goal, n_iter) == O.grow(goal, n_iter)
        assert_same_tree(G, O)
        w = G.wave_stats()
        rounds.append((w["rounds_parallel"] + w["rounds_sequential"] - G.tail_commits(), w["rounds_sequential"], G.tail_commits()))
        G.close()
    (r_fix, seq_fix, _), (r_tail, seq_tail, t_tail), (r_off, seq_off, t_off) = rounds
    assert t_off == 0 and t_tail > 0
    assert seq_fix < seq_tail // 2       # most of these samples are the simple parent-and-child case
    assert r_fix < r_tail < r_off


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct,n_iter", [(1, 0.1, 20000), (0, 0.0, 60000)])
def test_propagated_costs_are_forwarded(gpu_required, monkeypatch, dynamic, p_direct, n_iter):
    """A near node whose cost an earlier sample of the sub-round lowers by PROPAGATION (it lies below a node that sample
    rewires): the writer leaves the new cost on record and the reader re-checks its decision with it (KRRT_PROP_FORWARD)
    instead of ending the prefix.  Same tree and counters either way, fewer sub-rounds with it."""
    goal = scenes.GOAL if dynamic else None
    rounds = []
    for fwd in ("0", "1"):
        monkeypatch.setenv("KRRT_PROP_FORWARD", fwd)
        G, O = make_pair(gpu_required, dynamic=dynamic, p_direct=p_direct, capacity=n_iter + 2)
        assert G.grow(goal, n_iter) == O.grow(goal, n_iter)
        assert_same_tree(G, O)
        rounds.append(G.wave_stats()["rounds_parallel"])
        G.close()
    assert rounds[1] < rounds[0]


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct,n_iter", [(1, 0.1, 20000), (0, 0.0, 60000)])
def test_newcomer_edges_tested_at_validation(gpu_required, monkeypatch, dynamic, p_direct, n_iter):
    """A near neighbour inserted by an earlier sample of the same sub-round that would be tried as parent, or is a rewire
    candidate: its edge is collision-tested at validation (KRRT_JOIN_TEST); if it collides the decision stands and the
    test is counted, as a sequential planner counts it.  Same tree and counters either way, fewer sub-rounds with it."""
    goal = scenes.GOAL if dynamic else None
    rounds = []
    for on in ("0", "1"):
        monkeypatch.setenv("KRRT_JOIN_TEST", on)
        G, O = make_pair(gpu_required, dynamic=dynamic, p_direct=p_direct, capacity=n_iter + 2)
        assert G.grow(goal, n_iter) == O.grow(goal, n_iter)
        assert_same_tree(G, O)
        rounds.append(G.wave_stats()["rounds_parallel"])
        G.close()
    assert rounds[1] < rounds[0]


@pytest.mark.gpu
def test_dataflow_sub_rounds_same_tree(gpu_required, monkeypatch):
    """KRRT_FLOW=1 (experimental, off by default: measured no faster): inside a sub-round a sample waits for the samples
    before it only (per-sample flags with release/acquire) instead of two of the three grid barriers.  Same tree, same
    counters, same sub-rounds."""
    rounds = []
    for flow in ("0", "1"):
        monkeypatch.setenv("KRRT_FLOW", flow)
        G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, capacity=20002)
        assert G.grow(scenes.GOAL, 20000) == O.grow(scenes.GOAL, 20000)
        assert_same_tree(G, O)
        rounds.append(G.wave_stats())
        G.close()
    assert rounds[0] == rounds[1]


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_grow_headline_size_other_seeds(gpu_required, seed):
    """The headline plan (C2: 9 moving opponents, 20 000 iterations) on other sample streams: tree and counters equal the
    oracle's (the rare branches of the parallel commit — second cost test of a candidate child, chains, colliding candidates
    that stop being candidates — are hit by different streams)."""
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, capacity=20002, seed=seed)
    assert G.grow(scenes.GOAL, 20000) == O.grow(scenes.GOAL, 20000)
    assert_same_tree(G, O)
    G.close()


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct", [(0, 0.0), (1, 0.0), (1, 0.1)])
def test_goal_connection_rule(gpu_required, dynamic, p_direct):
    """O8 (KRRT_GOAL_CONNECT, the default): the goal node behind the tree, its parent, cost and arrival time, the chain
    and the fine trajectory through it — bit-identical with the oracle, at every call of a growing tree."""
    G, O = make_pair(gpu_required, dynamic=dynamic, p_direct=p_direct, capacity=4096)
    seen_solved = False
    for n_iter in (0, 1, 9, 15, 25, 50, 400, 3000):
        rg = G.grow(scenes.GOAL, n_iter)
        ro = O.grow(scenes.GOAL, n_iter)
        assert rg == ro
        assert_same_tree(G, O)
        if not ro[0]:
            assert rg[1] == -1
            continue
        seen_solved = True
        assert ro[1] > 0
        a, b = G.result(*rg), O.result(*ro)
        assert a.tobytes() == b.tobytes()
        ps, pt = G.path(rg[1])
        qs, qt = O.path(ro[1])
        assert np.array_equal(ps, qs) and np.array_equal(pt, qt) and np.array_equal(ps[-1], scenes.GOAL)
        ref = np.concatenate([ob.solve(qs[k], qs[k + 1], 0.5, 0.1, qt[k]) for k in range(len(qs) - 1)])
        assert np.array_equal(G.trajectory(rg[1], 0.1), ref)
        if ro[1] == O.size:   # the goal node's stored edge follows the tree's in tree.trajectories
            rows = np.empty((O.size + 1, 10, 5))
            G._ck(capi.lib().krrt_tree_edges(G.h, 10, capi._d(rows), O.size + 1))
            tau, _ = ob.steer(qs[-2], qs[-1])
            want = np.array([np.append(ob.edge_state(qs[-2], qs[-1], tau[0], k / 9.0), qt[-2] + tau[0] * (k / 9.0))
                             for k in range(10)])
            assert np.array_equal(rows[-1], want)
    assert seen_solved
    # without a goal nothing is solved and the goal node's index is no longer accepted
    assert G.grow(None, 5) == O.grow(None, 5) == (False, -1)
    with pytest.raises(capi.KrrtError):
        G.path(G.size)


@pytest.mark.gpu
def test_goal_tolerance_mode_and_full_tree(gpu_required):
    """KRRT_GOAL_TOLERANCE keeps the tolerance-ball rule; a FULL tree still has room for the connected goal node."""
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.1, capacity=4096, goal_mode=1, goal_tol=0.5)
    rg, ro = G.grow(scenes.GOAL, 3000), O.grow(scenes.GOAL, 3000)
    assert rg == ro and ro[0] and ro[1] < O.size
    assert G.result(*rg).tobytes() == O.result(*ro).tobytes()
    G, O = make_pair(gpu_required, dynamic=1, p_direct=0.0, capacity=300)
    rg, ro = G.grow(scenes.GOAL, 600), O.grow(scenes.GOAL, 600)
    assert rg == ro and ro[0] and ro[1] == 300 == O.size
    assert G.result(*rg).tobytes() == O.result(*ro).tobytes()
    # set_params switches the rule on a live context
    cfg = G.cfg
    cfg.goal_mode = 1
    G._ck(capi.lib().krrt_set_params(G.h, ctypes.byref(cfg)))
    assert G.grow(scenes.GOAL, 0) == (False, -1)


@pytest.mark.gpu
@pytest.mark.parametrize("goal_mode,dynamic,p_direct", [(0, 1, 0.0), (0, 0, 0.1), (1, 1, 0.1)])
def test_plan_batch_goal_rule(gpu_required, goal_mode, dynamic, p_direct):
    starts, goals, obs = scenes.batch_queries(400, 9, seed=23)
    goals[::7] = starts[::7]                      # start == goal
    goals[3, :2] = obs[3, 0, :2]                  # goal inside an obstacle: every offer collides at its end point
    kw = dict(p_direct=p_direct, seed=9, goal_mode=goal_mode, goal_tol=0.7)
    G = capi.Planner(capacity=1 << 12, **kw)
    for n_iter in (200, 30):
        out = G.plan_batch(starts, goals, obs, dynamic, n_iter, query_base=77)
        ref = ob.plan_batch(ob.default_config(capacity=1 << 12, **kw), starts, goals, obs, dynamic, n_iter, query_base=77,
                            nthreads=8)
        assert out.tobytes() == ref.tobytes()
        if goal_mode == 0:
            assert out["solved"].mean() > (0.9 if n_iter == 200 else 0.5)
            assert dynamic or not out["solved"][3]    # (a moving disc leaves the goal free sooner or later)
            conn = out["goal_index"] == out["n_nodes"]
            assert conn.sum() > 0 and np.all(out["goal_index"][out["solved"] == 1] > 0)
    # the v1 kernel (every list in shared memory) applies the same rule
    import os
    os.environ["KRRT_BATCH_V1"] = "1"
    try:
        G1 = capi.Planner(capacity=1 << 12, **kw)
        out1 = G1.plan_batch(starts, goals, obs, dynamic, 30, query_base=77)
    finally:
        del os.environ["KRRT_BATCH_V1"]
    assert out1.tobytes() == ref.tobytes()


@pytest.mark.parametrize("dynamic,p_direct,goal_mode,n_iter", [(1, 0.1, 0, 75), (0, 0.0, 0, 100), (1, 0.2, 1, 200), (1, 0.1, 0, 256)])
def test_grow_logged_equals_single_calls(gpu_required, dynamic, p_direct, goal_mode, n_iter):
    """krrt_grow_logged: ONE device call for the caller's `for(i<iteration) rrt.grow(xg)` loop (…node.cpp:176-185).  Record i is
    what single call i returns (solved, goalIndex) and what the loop reads after it (tree size, last sample); the tree after the
    call is the tree of the single calls.  Forced goal samples (…node.cpp:183-184) ride along as explicit rows."""
    kw = dict(goal_mode=goal_mode, goal_tol=0.8 if goal_mode else 0.0)
    G, O = make_pair(gpu_required, dynamic=dynamic, capacity=2048, p_direct=p_direct, **kw)
    rows = np.full((n_iter, 4), np.nan)
    rows[n_iter // 2 + 1] = scenes.GOAL
    rows[n_iter - 1] = scenes.GOAL
    rec, solved, gi = G.grow_logged(scenes.GOAL, n_iter, rows)
    n_prev = 1
    for i in range(n_iter):
        so, go = O.grow(scenes.GOAL, 1, samples=rows[i:i + 1])
        assert (bool(rec["solved"][i]), int(rec["goal_index"][i])) == (so, go), i
        assert rec["n_nodes"][i] == O.size and bool(rec["inserted"][i]) == (O.size > n_prev), i
        assert np.array_equal(rec["sample"][i], O.last_sample()), i
        if so:
            assert rec["goal_cost"][i] == O.result(so, go)["cost"], i
        n_prev = O.size
    assert (solved, gi) == (so, go)
    assert_same_tree(G, O)
    # a second logged call continues the same tree
    rec2, solved2, gi2 = G.grow_logged(scenes.GOAL, 40)
    for i in range(40):
        so, go = O.grow(scenes.GOAL, 1)
        assert (bool(rec2["solved"][i]), int(rec2["goal_index"][i]), int(rec2["n_nodes"][i])) == (so, go, O.size), i
    assert_same_tree(G, O)
    with pytest.raises(capi.KrrtError) as e:
        G.grow_logged(scenes.GOAL, 257)
    assert e.value.code == capi.KRRT_ERR_UNSUPPORTED


def test_grow_logged_edge_cases(gpu_required):
    """One iteration, no goal, a full tree, explicit rows only, and a logged call after plain ones: records against single calls."""
    G, O = make_pair(gpu_required, dynamic=1, capacity=40, p_direct=0.3)
    G.grow(scenes.GOAL, 7)
    O.grow(scenes.GOAL, 7)
    rec, solved, gi = G.grow_logged(None, 1)                    # no goal: never solved
    so, go = O.grow(None, 1)
    assert (bool(rec["solved"][0]), int(rec["goal_index"][0]), int(rec["n_nodes"][0])) == (so, go, O.size) == (False, -1, O.size)
    rows = np.array([[1.0, 2.0, 0.5, -0.5], [1.0, 2.0, 0.5, -0.5], [-3.0, 1.0, 0.0, 0.25]])   # the second one is a duplicate
    rec, solved, gi = G.grow_logged(scenes.GOAL, 3, rows)
    for i in range(3):
        so, go = O.grow(scenes.GOAL, 1, samples=rows[i:i + 1])
        assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i])) == (so, go, O.size), i
        assert np.array_equal(rec["sample"][i], rows[i])
    assert not rec["inserted"][1]
    rec, solved, gi = G.grow_logged(scenes.GOAL, 120)           # the tree fills up (capacity 40) on the way
    n_prev = O.size
    for i in range(120):
        so, go = O.grow(scenes.GOAL, 1)
        if O.size == 40 and n_prev == 40:
            break                                                # from here on single calls find the tree full
        assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i])) == (so, go, O.size), i
        n_prev = O.size
    assert G.size == 40 and G.grow_done()[1]
    # iteration i is the one that found the tree full: it ends the call, and it is a record of its own (so, go: the full tree's)
    assert 0 < i < 119 and G.grow_done()[0] == i + 1
    assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i]), int(rec["inserted"][i])) == (so, go, 40, 0)
    # the iterations behind it were never run: each repeats that record (include/krrt.h), none shows stale or unwritten log entries
    for j in range(i + 1, 120):
        for f in ("n_nodes", "solved", "goal_index", "inserted", "goal_cost"):
            assert rec[f][j] == rec[f][i], (j, f)
        assert np.array_equal(rec["sample"][j], rec["sample"][i]), j
    # a second logged call on the full tree: its first iteration finds the tree full, the rest repeat it
    rec2, solved2, gi2 = G.grow_logged(scenes.GOAL, 9)
    so, go = O.grow(scenes.GOAL, 1)
    assert G.grow_done()[0] == 1 and (bool(solved2), gi2) == (so, go)
    for j in range(9):
        assert (bool(rec2["solved"][j]), int(rec2["goal_index"][j]), int(rec2["n_nodes"][j]), int(rec2["inserted"][j])) == (so, go, 40, 0), j
        assert np.array_equal(rec2["sample"][j], rec2["sample"][0])
