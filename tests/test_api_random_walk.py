"""Differential random walk over the single-tree C ABI: a seeded sequence of calls — grow in pieces of every size (with and
without a goal, with explicit sample rows and NaN rows), logged grows, restarts, a new obstacle set in the middle of a plan,
radius queries, paths, result records — is issued to a context of libkrrt and to the CPU checker of the same precision;
after every call that returns something the two must agree exactly, and every few calls the whole tree and all ten counters
are compared.  The single calls are covered one by one elsewhere; this test is about their ORDER (state left behind by one
call and read by the next: the goal node's slot behind the tree, the grid index across restarts, the logged call's buffers,
the obstacle set of a running plan)."""
import numpy as np
import pytest

import oracle_binding as ob
import oracle_f32_binding as of
from robosoccer_motion_planning_b200 import capi, scenes

pytestmark = pytest.mark.gpu

CAP = 6000


def _pair(prec, seed):
    kw = dict(capacity=CAP, p_direct=0.15, seed=seed)
    G = capi.Planner(precision=prec, **kw)
    O = of.OraclePlannerF32(**kw) if prec == capi.KRRT_F32 else ob.OraclePlanner(**kw)
    return G, O


def _same_tree(G, O, tag):
    assert G.size == O.size and G.iteration == O.iteration, tag
    for a, b, what in zip(G.tree(), O.tree(), ("states", "parents", "cost-to-come", "times")):
        assert np.array_equal(a, b), (tag, what)
    assert G.stats() == O.stats(), tag


def _random_obstacles(rng, m, dynamic):
    obs = np.zeros((m, 4))
    obs[:, 0] = rng.uniform(scenes.FIELD_LO[0], scenes.FIELD_HI[0], m)
    obs[:, 1] = rng.uniform(scenes.FIELD_LO[1], scenes.FIELD_HI[1], m)
    if dynamic:
        obs[:, 2:] = rng.uniform(-1.0, 1.0, (m, 2))
    # keep the start region free so that a plan can leave it
    keep = np.hypot(obs[:, 0] - scenes.START[0], obs[:, 1] - scenes.START[1]) > 1.0
    return obs[keep]


@pytest.mark.parametrize("prec", [capi.KRRT_F64, capi.KRRT_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("seed", list(range(1, 9)))
def test_random_call_sequences_agree_with_the_checker(gpu_required, prec, seed):
    rng = np.random.default_rng(1000 * seed + (7 if prec == capi.KRRT_F32 else 0))
    G, O = _pair(prec, seed)
    dynamic = int(rng.integers(0, 2))
    obs = scenes.dynamic_obstacles() if dynamic else scenes.static_obstacles()
    for P in (G, O):
        P.set_obstacles(obs, dynamic)
        P.set_start(scenes.START)
    goal = np.array(scenes.GOAL, dtype=float)
    last = None  # (solved, goal_index) of the last grow, valid until the tree changes
    ops = ["grow", "grow", "grow", "grow_rows", "logged", "near", "path", "result", "restart", "obstacles", "goal", "tree"]
    for step in range(140):
        op = ops[int(rng.integers(0, len(ops)))]
        tag = (step, op)
        if op in ("grow", "grow_rows"):
            n = int(rng.choice([0, 1, 2, 5, 31, 64, 200, 700]))
            g = goal if rng.random() < 0.8 else None
            rows = None
            if op == "grow_rows" and n > 0:
                rows = np.full((n, 4), np.nan)
                k = rng.random(n) < 0.3
                rows[k] = rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, (int(k.sum()), 4))
                if n > 1:
                    rows[n - 1] = goal
            rg = G.grow(g, n, samples=rows)
            ro = O.grow(g, n, samples=rows)
            assert rg == ro, tag
            assert G.size == O.size and G.iteration == O.iteration, tag
            assert np.array_equal(G.last_sample(), O.last_sample()) or G.iteration == 0, tag
            last = rg if g is not None else None
        elif op == "logged":
            n = int(rng.choice([1, 7, 75, 256]))
            if G.size + n > 1024:
                with pytest.raises(capi.KrrtError) as e:
                    G.grow_logged(goal, n)
                assert e.value.code == capi.KRRT_ERR_UNSUPPORTED, tag
                continue
            rec, sg, gg = G.grow_logged(goal, n)
            ro = O.grow(goal, n)
            assert (sg, gg) == ro, tag
            assert int(rec["n_nodes"][-1]) == O.size and np.array_equal(rec["sample"][-1], O.last_sample()), tag
            last = ro
        elif op == "near":
            x = rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, 4)
            r2 = float(rng.choice([0.0, 0.5, 4.0, 100.0]))
            gi, gn, gd = G.near(x, r2, cap=CAP)
            oi, on, od = O.near(x, r2, cap=CAP)
            assert np.array_equal(gi, oi) and gn == on and gd == od, tag
        elif op == "path":
            idx = last[1] if (last is not None and last[0]) else int(rng.integers(0, G.size))
            gs, gt = G.path(idx)
            os_, ot = O.path(idx)
            assert np.array_equal(gs, os_) and np.array_equal(gt, ot), tag
        elif op == "result":
            if last is None:
                continue
            rg, ro = G.result(*last), O.result(*last)
            assert rg.tobytes() == ro.tobytes(), tag
        elif op == "restart":
            x0 = np.array(scenes.START, dtype=float)
            if rng.random() < 0.5:
                x0[:2] += rng.uniform(-0.3, 0.3, 2)
            for P in (G, O):
                P.set_start(x0)
            last = None
            assert G.size == 1 and O.size == 1, tag
        elif op == "obstacles":
            dynamic = int(rng.integers(0, 2))
            m = int(rng.integers(0, 12))
            obs = _random_obstacles(rng, m, dynamic)
            for P in (G, O):
                P.set_obstacles(obs, dynamic)
            last = None
        elif op == "goal":
            goal = np.array(scenes.GOAL, dtype=float)
            if rng.random() < 0.7:
                goal[:2] += rng.uniform(-1.0, 1.0, 2)
            last = None
        else:
            _same_tree(G, O, tag)
        if G.size > CAP - 800:  # keep clear of a full tree (covered by test_capacity_full): start again
            for P in (G, O):
                P.set_start(scenes.START)
            last = None
    _same_tree(G, O, "end")
    G.close()
    O.close()
