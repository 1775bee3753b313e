"""fp32 mode (KRRT_F32 contexts; BASELINE.json north_star: "1e-4 in fp32 mode").

The binary32 planner is the binary64 planner's source compiled in float on both sides (kernels: csrc/krrt_f32.cu; checker:
oracle/_build/libkrrt_oracle_f32.so), so GPU results compare BIT FOR BIT with the float checker; the float checker in turn
stays within north_star's 1e-4 of the binary64 one on costs and arrival times of identical edges (the tolerance is written
below), and plans with the same statistics."""
import numpy as np
import pytest

import oracle_binding as ob
import oracle_f32_binding as of

RTOL_F32 = 1e-4   # north_star: path costs and arrival times within 1e-4 relative in fp32 mode


def _pairs(n, seed=0):
    rng = np.random.default_rng(seed)
    lo = np.array([-9.0, -6.0, -2.0, -2.0])
    hi = np.array([9.0, 6.0, 2.0, 2.0])
    x0 = rng.uniform(lo, hi, (n, 4)).astype(np.float32).astype(np.float64)   # float-representable: both builds see the same edge
    x1 = rng.uniform(lo, hi, (n, 4)).astype(np.float32).astype(np.float64)
    return x0, x1


def test_float_checker_steering_within_1e4_of_double():
    x0, x1 = _pairs(200000)
    t32, c32 = of.steer(x0, x1)
    t64 = np.empty(len(x0))
    c64 = np.empty(len(x0))
    ob.lib().ko_steer_batch(ob._d(x0), ob._d(x1), len(x0), 0.5, ob._d(t64), ob._d(c64))
    rc = np.abs(c32 - c64) / c64
    rt = np.abs(t32 - t64) / t64
    assert rc.max() <= RTOL_F32, rc.max()
    # the arrival time of the optimum is as well conditioned except where two roots of the quartic cost the same to within
    # float rounding (three-root zone, SURVEY.md 8-M: 0.07 % of field-scale pairs) and the other one is picked
    assert np.mean(rt <= RTOL_F32) >= 0.999, np.mean(rt <= RTOL_F32)
    assert np.median(rc) < 1e-6


def test_float_checker_sampler_and_radius_follow_the_double_ones():
    c64 = ob.default_config(p_direct=0.1, query_id=7)
    c32 = of.default_config(p_direct=0.1, query_id=7)
    goal = np.array([7.0, 0.0, 0.0, 0.0])
    for it in range(200):
        a = ob.sample(c64, it, goal)
        b = of.sample(c32, it, goal)
        assert np.allclose(a, b, rtol=0, atol=2e-6)                 # same Philox words, rounded to binary32
    for n in (2, 10, 200, 5000):
        assert abs(of.rho2(c32, n) - ob.lib().ko_rho2(c64, n)) <= 1e-5 * ob.lib().ko_rho2(c64, n)


def test_float_checker_plans_like_the_double_one():
    from robosoccer_motion_planning_b200 import scenes
    starts, goals, obs = scenes.batch_queries(256, 9, seed=3)
    r64 = ob.plan_batch(ob.default_config(p_direct=0.1), starts, goals, obs, 1, 200, nthreads=8)
    r32 = of.plan_batch(of.default_config(p_direct=0.1), starts, goals, obs, 1, 200, nthreads=8)
    assert abs(r64["solved"].mean() - r32["solved"].mean()) <= 0.02
    both = (r64["solved"] != 0) & (r32["solved"] != 0)
    assert both.mean() > 0.9
    # trees diverge once a float rounding flips a decision (any RRT* does under perturbation); the plans' quality does not
    assert abs(r64["cost"][both].mean() - r32["cost"][both].mean()) <= 0.03 * r64["cost"][both].mean()
    # a plan's own record is consistent: arrival time of the last waypoint == reported time, first waypoint == start
    for rec, s in zip(r32[both][:32], starts[both][:32]):
        n = min(int(rec["path_len"]), 32)
        assert np.array_equal(rec["waypoints"][0, :4], s.astype(np.float32).astype(np.float64))
        if rec["path_len"] <= 32:
            assert rec["waypoints"][n - 1, 4] == rec["time"]


def test_float_checker_meets_north_star_against_the_double_one():
    """north_star's fp32 criterion, checker against checker: same parents / near sets (here: same tree sizes, goal nodes and
    path lengths) wherever no decision sat within float rounding of a tie, costs and arrival times within 1e-4."""
    from robosoccer_motion_planning_b200 import scenes
    starts, goals, obs = scenes.batch_queries(1024, 9, seed=0)
    r64 = ob.plan_batch(ob.default_config(p_direct=0.1), starts, goals, obs, 1, 200, nthreads=8)
    r32 = of.plan_batch(of.default_config(p_direct=0.1), starts, goals, obs, 1, 200, nthreads=8)
    same = (r64["n_nodes"] == r32["n_nodes"]) & (r64["goal_index"] == r32["goal_index"]) & (r64["path_len"] == r32["path_len"])
    assert same.mean() >= 0.99, same.mean()          # measured: 1024 of 1024
    ok = same & (r64["solved"] != 0)
    assert (np.abs(r32["cost"][ok] - r64["cost"][ok]) <= RTOL_F32 * r64["cost"][ok]).all()
    assert (np.abs(r32["time"][ok] - r64["time"][ok]) <= RTOL_F32 * r64["time"][ok]).all()
    wp = np.abs(r32["waypoints"][ok] - r64["waypoints"][ok])
    assert wp.max() <= 1e-4 * 18.0                  # waypoints: 1e-4 of the field's length


# ------------------------------------------------------------------------------------------------ GPU: bit-exact
@pytest.mark.gpu
def test_f32_steer_and_collide_bit_exact(gpu_required):
    from robosoccer_motion_planning_b200 import capi, scenes
    G = capi.Planner(precision=capi.KRRT_F32)
    x0, x1 = _pairs(100000, seed=5)
    t, c = G.steer_cost(x0, x1)
    to, co = of.steer(x0, x1)
    assert np.array_equal(t, to) and np.array_equal(c, co)
    # inputs that are NOT float-representable are rounded on load, exactly as the checker's binding rounds them
    y0, y1 = x0[:1000] + 1e-9, x1[:1000] - 1e-9
    t2, c2 = G.steer_cost(y0, y1)
    to2, co2 = of.steer(y0, y1)
    assert np.array_equal(t2, to2) and np.array_equal(c2, co2)
    for dynamic in (0, 1):
        obs = scenes.dynamic_obstacles()
        G.set_obstacles(obs, dynamic)
        n = 4000
        t0 = np.random.default_rng(1).uniform(0, 5, n)
        hit = G.edge_collide(x0[:n], x1[:n], t[:n], t0)
        ref = np.array([of.edge_collide(x0[i], x1[i], t[i], t0[i], obs, dynamic) for i in range(n)], dtype=np.uint8)
        assert np.array_equal(hit, ref)
        assert 0 < hit.sum() < n
    G.close()


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct,n_iter,goal_mode", [(1, 0.1, 200, 0), (0, 0.0, 75, 0), (1, 0.1, 120, 1)])
def test_f32_plan_batch_bit_exact(gpu_required, dynamic, p_direct, n_iter, goal_mode):
    from robosoccer_motion_planning_b200 import capi, scenes
    q = 512
    starts, goals, obs = scenes.batch_queries(q, 9, seed=11)
    kw = dict(p_direct=p_direct, goal_mode=goal_mode, goal_tol=0.75 if goal_mode else 0.0)
    G = capi.Planner(precision=capi.KRRT_F32, **kw)
    out = G.plan_batch(starts, goals, obs, dynamic, n_iter, query_base=100)
    ref = of.plan_batch(of.default_config(**kw), starts, goals, obs, dynamic, n_iter, query_base=100, nthreads=8)
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
    assert out["solved"].mean() > (0.9 if goal_mode == 0 else 0.05)
    # ... and it is not the binary64 plan: costs differ in the low bits
    G64 = capi.Planner(**kw)
    out64 = G64.plan_batch(starts, goals, obs, dynamic, n_iter, query_base=100)
    assert not np.array_equal(out["cost"], out64["cost"])
    both = (out["solved"] != 0) & (out64["solved"] != 0)
    assert abs(out["cost"][both].mean() - out64["cost"][both].mean()) <= 0.03 * out64["cost"][both].mean()
    G.close()
    G64.close()


@pytest.mark.gpu
def test_f32_connector_bit_exact(gpu_required):
    """Connector slot on a KRRT_F32 context: solver.solve rows, control inputs and the S stored samples of an edge."""
    from robosoccer_motion_planning_b200 import capi
    G = capi.Planner(precision=capi.KRRT_F32)
    x0, x1 = _pairs(300, seed=8)
    t, c = of.steer(x0, x1)
    for i in range(0, 300, 7):
        rows = G.solve(x0[i], x1[i], dt=0.1, t0=1.5)
        ref = of.solve(x0[i], x1[i], 0.5, 0.1, 1.5)
        assert rows.shape == ref.shape and np.array_equal(rows, ref), i
        rows2, inp = G.solve_inputs(x0[i], x1[i], dt=0.1)
        assert np.array_equal(inp, of.solve_inputs(x0[i], x1[i], 0.5, 0.1)), i
        assert np.array_equal(rows2, of.solve(x0[i], x1[i], 0.5, 0.1, 0.0))
    S = 10
    t0 = np.linspace(0.0, 3.0, 300)
    E = G.edge_samples(x0, x1, t0, S)
    for i in range(0, 300, 11):
        for k in range(S):
            h = np.float32(k) / np.float32(S - 1)
            st = of.edge_state(x0[i], x1[i], t[i], h)
            assert np.array_equal(E[i, k, :4], st), (i, k)
            assert E[i, k, 4] == float(np.float32(t0[i]) + np.float32(t[i]) * h)
    G.close()


@pytest.mark.gpu
def test_f32_near_overflow_takes_the_redo_pass(gpu_required):
    """A near radius that covers the whole field: every near set outgrows k_plan_batch2's 64 registers, the float k_plan_batch
    plans those queries, the result is the checker's."""
    from robosoccer_motion_planning_b200 import capi, scenes
    starts, goals, obs = scenes.batch_queries(64, 9, seed=2)
    kw = dict(p_direct=0.1, gamma=40.0, rho_max=30.0)
    G = capi.Planner(precision=capi.KRRT_F32, **kw)
    out = G.plan_batch(starts, goals, obs, 1, 150)
    ref = of.plan_batch(of.default_config(**kw), starts, goals, obs, 1, 150, nthreads=8)
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
    G.close()


def _pair32(dynamic, capacity, **kw):
    from robosoccer_motion_planning_b200 import capi, scenes
    G = capi.Planner(precision=capi.KRRT_F32, capacity=capacity, **kw)
    O = of.OraclePlannerF32(capacity=capacity, **kw)
    obs = scenes.dynamic_obstacles() if dynamic else scenes.static_obstacles()
    for P in (G, O):
        P.set_obstacles(obs, dynamic)
        P.set_start(scenes.START)
    return G, O


def _same_tree32(G, O):
    assert G.size == O.size and G.iteration == O.iteration
    for a, b, what in zip(G.tree(), O.tree(), ("states", "parents", "cost-to-come", "times")):
        assert np.array_equal(a, b), what
    assert G.stats() == O.stats()


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic,p_direct,n_iter,goal_mode", [(1, 0.1, 8000, 0), (0, 0.0, 6000, 0), (1, 0.2, 3000, 1)])
def test_f32_single_tree_bit_exact(gpu_required, dynamic, p_direct, n_iter, goal_mode):
    """The planner proper on a KRRT_F32 context: k_grow compiled in float against the float checker — states, parents, costs,
    times, all ten counters, goal node, path and result record; in several calls, with explicit rows, after a restart."""
    from robosoccer_motion_planning_b200 import scenes
    kw = dict(p_direct=p_direct, goal_mode=goal_mode, goal_tol=1.0 if goal_mode else 0.0)
    G, O = _pair32(dynamic, n_iter + 2, **kw)
    rows = np.full((100, 4), np.nan)
    rows[7] = scenes.GOAL
    rows[50] = [1.0 + 1e-9, -2.0, 0.5, 0.25]            # (not float-representable: rounded on the way in, on both sides)
    rg = G.grow(scenes.GOAL, 100, rows)
    ro = O.grow(scenes.GOAL, 100, rows)
    assert rg == ro
    assert np.array_equal(G.last_sample(), O.last_sample())
    rg = G.grow(scenes.GOAL, n_iter - 100)
    ro = O.grow(scenes.GOAL, n_iter - 100)
    assert rg == ro
    _same_tree32(G, O)
    if rg[0]:
        ps, pt = G.path(rg[1])
        qs, qt = O.path(ro[1])
        assert np.array_equal(ps, qs) and np.array_equal(pt, qt)
        assert G.result(*rg).tobytes() == O.result(*ro).tobytes()
    for P in (G, O):                                       # restart on the same context
        P.set_start(scenes.GOAL)
        P.grow(scenes.START, 500)
    _same_tree32(G, O)
    G.close()


@pytest.mark.gpu
def test_f32_single_tree_large_and_full(gpu_required):
    """30 000 iterations without a goal (grid index from 8192 nodes on, windows of 1024 samples), then a tree that fills up."""
    G, O = _pair32(0, 30002)
    G.grow(None, 30000)
    O.grow(None, 30000)
    _same_tree32(G, O)
    assert G.size > 20000
    G.close()
    from robosoccer_motion_planning_b200 import scenes
    G, O = _pair32(1, 300, p_direct=0.1)
    for P in (G, O):
        P.grow(scenes.GOAL, 600)
    assert G.size == O.size == 300 and G.grow_done()[1]
    _same_tree32(G, O)
    G.close()


@pytest.mark.gpu
def test_f32_tree_close_to_the_binary64_tree(gpu_required):
    """north_star's fp32 criterion on a single tree: the float tree has the binary64 tree's parents wherever no decision sat within
    float rounding of a tie — here for the first few hundred expansions of a plan — and costs / times within 1e-4 there."""
    from robosoccer_motion_planning_b200 import capi, scenes
    G32 = capi.Planner(precision=capi.KRRT_F32, capacity=4096, p_direct=0.1)
    G64 = capi.Planner(capacity=4096, p_direct=0.1)
    for P in (G32, G64):
        P.set_obstacles(scenes.dynamic_obstacles(), 1)
        P.set_start(scenes.START)
        P.grow(scenes.GOAL, 400)
    s32, p32, c32, t32 = G32.tree()
    s64, p64, c64, t64 = G64.tree()
    n = min(len(p32), len(p64))
    same = p32[:n] == p64[:n]
    assert len(p32) == len(p64) and same.mean() >= 0.99
    ok = same & (c64[:n] > 0)
    assert (np.abs(c32[:n][ok] - c64[:n][ok]) <= RTOL_F32 * c64[:n][ok]).all()
    assert (np.abs(t32[:n][ok] - t64[:n][ok]) <= RTOL_F32 * t64[:n][ok]).all()
    G32.close()
    G64.close()


@pytest.mark.gpu
def test_f32_grow_logged_equals_single_calls(gpu_required):
    """krrt_grow_logged on a binary32 tree: record i is what single call i returns and what the caller's loop reads after it."""
    from robosoccer_motion_planning_b200 import scenes
    G, O = _pair32(1, 2048, p_direct=0.1)
    n_iter = 200
    rows = np.full((n_iter, 4), np.nan)
    rows[101] = scenes.GOAL
    rows[199] = scenes.GOAL
    rec, solved, gi = G.grow_logged(scenes.GOAL, n_iter, rows)
    n_prev = 1
    for i in range(n_iter):
        so, go = O.grow(scenes.GOAL, 1, samples=rows[i:i + 1])
        assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i])) == (so, go, O.size), i
        assert bool(rec["inserted"][i]) == (O.size > n_prev) and np.array_equal(rec["sample"][i], O.last_sample()), i
        if so:
            assert rec["goal_cost"][i] == O.result(so, go)["cost"], i
        n_prev = O.size
    assert (solved, gi) == (so, go)
    _same_tree32(G, O)
    G.close()


@pytest.mark.gpu
def test_f32_grow_logged_on_a_tree_that_fills_up(gpu_required):
    """binary32: the iteration that finds the tree full ends the logged call with a record of its own; the iterations
    behind it repeat that record (include/krrt.h) — no stale or unwritten log entry shows through"""
    from robosoccer_motion_planning_b200 import scenes
    G, O = _pair32(1, 40, p_direct=0.3)
    rec, solved, gi = G.grow_logged(scenes.GOAL, 120)
    n_prev = O.size
    for i in range(120):
        so, go = O.grow(scenes.GOAL, 1)
        if O.size == 40 and n_prev == 40:
            break
        assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i])) == (so, go, O.size), i
        n_prev = O.size
    assert G.size == 40 and G.grow_done() == (i + 1, True) and 0 < i < 119
    assert (bool(rec["solved"][i]), int(rec["goal_index"][i]), int(rec["n_nodes"][i]), int(rec["inserted"][i])) == (so, go, 40, 0)
    for j in range(i + 1, 120):
        for f in ("n_nodes", "solved", "goal_index", "inserted", "goal_cost"):
            assert rec[f][j] == rec[f][i], (j, f)
        assert np.array_equal(rec["sample"][j], rec["sample"][i]), j
    G.close()


@pytest.mark.gpu
def test_f32_tree_edges_and_trajectory(gpu_required):
    """tree.trajectories (robosoccer_visual.hpp:18) and the fine re-solve of the found path (…node.cpp:232-249) on a binary32 tree."""
    from robosoccer_motion_planning_b200 import scenes
    G, O = _pair32(1, 2002, p_direct=0.1)
    rg = G.grow(scenes.GOAL, 1500)
    ro = O.grow(scenes.GOAL, 1500)
    assert rg == ro and rg[0]
    st, par, ctc, tim = O.tree()
    S = 10
    E = G.tree_edges(S)
    n = len(par) + (1 if rg[1] == len(par) else 0)
    assert E.shape[0] == n
    assert np.array_equal(E[0, :, :4], np.broadcast_to(st[0], (S, 4))) and not E[0, :, 4].any()
    tau, _ = of.steer(st[par[1:]], st[1:])
    for i in range(1, len(par), 37):
        for k in (0, 3, S - 1):
            h = np.float32(k) / np.float32(S - 1)
            assert np.array_equal(E[i, k, :4], of.edge_state(st[par[i]], st[i], tau[i - 1], h)), (i, k)
            assert E[i, k, 4] == float(np.float32(tim[par[i]]) + np.float32(tau[i - 1]) * h)
    rows = G.trajectory(rg[1], dt=0.1)
    path, ptime = O.path(ro[1])
    want = np.concatenate([of.solve(a, b, 0.5, 0.1, t0) for a, b, t0 in zip(path[:-1], path[1:], ptime[:-1])])
    assert rows.shape == want.shape and np.array_equal(rows, want)
    G.close()


@pytest.mark.gpu
def test_f32_near_query(gpu_required):
    """krrt_near on a binary32 tree: indices, nearest node and its squared distance are the float checker's."""
    from robosoccer_motion_planning_b200 import scenes
    G, O = _pair32(0, 8192)
    l0 = G.launch_count
    for P in (G, O):
        P.grow(None, 4000)
    # krrt_launch_count on a binary32 tree is the tree's own counter: k_begin_grow, k_grow and the goal kernel of this call
    assert G.launch_count - l0 == 3
    rng = np.random.default_rng(4)
    for k in range(20):
        x = rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, 4)
        r2 = [0.0, 0.3, 1.0, 9.0, 400.0][k % 5]
        cap = 8192 if k % 2 == 0 else 64                  # (the query's device buffer is kept between calls and only grows)
        gi, gn, gd = G.near(x, r2, cap=cap)
        oi, on, od = O.near(x, r2, cap=8192)
        assert np.array_equal(gi, oi[:cap]) and gn == on and gd == od, k
    assert G.launch_count - l0 == 3 + 20                  # one k_near_query per query
    s = O.tree()[0][123]
    gi, gn, gd = G.near(s, 0.0)
    assert list(gi) == [123] and gn == 123 and gd == 0.0
    G.close()


@pytest.mark.gpu
def test_f32_edge_cases(gpu_required):
    """Empty and ragged inputs on a KRRT_F32 context: no queries, one query, one iteration, no obstacles, a start that coincides
    with the goal, the largest obstacle count — every record equals the float checker's."""
    from robosoccer_motion_planning_b200 import capi, scenes
    G = capi.Planner(precision=capi.KRRT_F32, p_direct=0.25)
    cfg = of.default_config(p_direct=0.25)
    starts, goals, obs = scenes.batch_queries(33, 9, seed=21)
    assert len(G.plan_batch(starts[:0], goals[:0], obs[:0], 1, 50)) == 0
    for q, m, n_iter, dyn in ((1, 9, 1, 1), (1, 9, 200, 0), (33, 0, 40, 1), (7, 3, 2, 1)):
        o = obs[:q, :m]
        out = G.plan_batch(starts[:q], goals[:q], o, dyn, n_iter, query_base=5)
        ref = of.plan_batch(cfg, starts[:q], goals[:q], o, dyn, n_iter, query_base=5, nthreads=4)
        assert np.array_equal(out.view(np.uint8), ref.view(np.uint8)), (q, m, n_iter, dyn)
    same = goals[:4].copy()
    out = G.plan_batch(same, same, obs[:4], 1, 30)              # start == goal: every sample that is the goal is a duplicate of the root
    ref = of.plan_batch(cfg, same, same, obs[:4], 1, 30, nthreads=2)
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
    rng = np.random.default_rng(3)
    big = np.concatenate([rng.uniform(-9, 9, (2, 64, 1)), rng.uniform(-6, 6, (2, 64, 1)), rng.uniform(-1, 1, (2, 64, 2))], axis=2)
    big[:, :, :2] += 30.0                                       # 64 discs (KRRT_MAX_OBS) parked off the field
    out = G.plan_batch(starts[:2], goals[:2], big, 1, 60)
    ref = of.plan_batch(cfg, starts[:2], goals[:2], big, 1, 60, nthreads=2)
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
    assert G.steer_cost(starts[:0], goals[:0])[0].shape == (0,)
    G.close()
