"""CPU tests of the oracle itself (parity is UNPINNED against the reference — the RRT* sources are an absent
submodule — so the oracle is pinned by the published math instead: SURVEY.md §8-M identities, brute-force
minimisation, the Random123 Philox known-answer vectors, tree invariants, committed golden vectors)."""
import ctypes
import os

import numpy as np
import pytest

import oracle_binding as ob
from robosoccer_motion_planning_b200 import scenes

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def rand_states(rng, n):
    return rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, (n, 4))


def cost_of_tau(a, b, r, tau):
    dp = b[:2] - a[:2]
    v0, v1 = a[2:], b[2:]
    return tau + r * (12 * dp @ dp / tau**3 - 12 * dp @ (v0 + v1) / tau**2 + 4 * (v0 @ v0 + v0 @ v1 + v1 @ v1) / tau)


def test_philox_known_answers():
    """Random123 kat_vectors, philox4x32-10."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
        ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
        ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
         (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
    ]
    out = (ctypes.c_uint32 * 4)()
    for ctr, key, want in kat:
        ob.lib().ko_philox(key[0] | (key[1] << 32), *ctr, out)
        assert tuple(out) == want


def test_steer_identities():
    rng = np.random.default_rng(0)
    n = 20000
    a, b = rand_states(rng, n), rand_states(rng, n)
    r = 0.5
    tau, c = ob.steer(a, b, r)
    assert np.all(tau > 0) and np.all(np.isfinite(c))
    dp = b[:, :2] - a[:, :2]
    s = a[:, 2:] + b[:, 2:]
    A = 4 * r * ((a[:, 2:] ** 2).sum(1) + (a[:, 2:] * b[:, 2:]).sum(1) + (b[:, 2:] ** 2).sum(1))
    Bq = 24 * r * (dp * s).sum(1)
    C0 = 36 * r * (dp ** 2).sum(1)
    res = tau**4 - A * tau**2 + Bq * tau - C0
    assert np.max(np.abs(res) / (tau**4 + A * tau**2 + np.abs(Bq) * tau + C0)) < 1e-12   # quartic residual
    cc = tau + A / tau - 0.5 * Bq / tau**2 + C0 / 3 / tau**3
    assert np.allclose(c, cc, rtol=1e-12)
    for i in range(0, n, 400):   # end point hit + cost == integral of (1 + r u^2) by quadrature
        e = ob.edge_state(a[i], b[i], tau[i], 1.0)
        assert np.allclose(e, b[i], atol=1e-9)
        h = np.linspace(0, 1, 2001)
        st = np.array([ob.edge_state(a[i], b[i], tau[i], x) for x in h])
        u = np.gradient(st[:, 2:], h * tau[i], axis=0, edge_order=2)
        J = np.trapezoid(1 + r * (u ** 2).sum(1), h * tau[i])
        assert abs(J - c[i]) < 2e-4 * c[i]


def test_steer_is_global_minimum():
    """all-roots rule: c* <= c(tau) on a dense grid, including the 3-root zone (fast and close states)."""
    rng = np.random.default_rng(1)
    a, b = rand_states(rng, 3000), rand_states(rng, 3000)
    a[::2, 2:] *= 4; b[::2, 2:] *= 4
    b[::2, :2] = a[::2, :2] + 0.3 * rng.uniform(-1, 1, (1500, 2))
    tau, c = ob.steer(a, b, 0.5)
    grid = np.geomspace(1e-3, 60, 6000)
    for i in range(3000):
        cg = np.array(cost_of_tau(a[i], b[i], 0.5, grid))
        assert c[i] <= cg.min() * (1 + 1e-9)
        assert abs(cost_of_tau(a[i], b[i], 0.5, tau[i]) - c[i]) <= 1e-9 * c[i]


def test_steer_degenerate():
    z = np.array([1.0, 2.0, 0.0, 0.0])
    t, c = ob.steer(z, z)
    assert t[0] == 0.0 and c[0] == 0.0
    t, c = ob.steer([0, 0, 1, 0], [0, 0, -1, 0])          # same position, reverse velocity
    assert t[0] > 0 and np.isfinite(c[0])
    t, c = ob.steer([0, 0, 0, 0], [4, 0, 0, 0], r=0.5)    # rest-to-rest: tau^4 = 36 r d^2
    assert np.isclose(t[0], (36 * 0.5 * 16) ** 0.25, rtol=1e-14)


def test_solve_rows():
    a, b = np.array([0.0, 0, 1, 0]), np.array([3.0, 1, 0, 1])
    rows = ob.solve(a, b, 0.5, 0.1, 2.0)
    tau, _ = ob.steer(a, b)
    assert np.array_equal(rows[0, :4], a) and rows[0, 4] == 2.0
    assert np.allclose(rows[-1, :4], b, atol=1e-12) and rows[-1, 4] == 2.0 + tau[0]
    assert len(rows) == int(np.ceil(tau[0] / 0.1)) + 1
    assert np.all(np.diff(rows[:, 4]) > 0)


def test_edge_collide_static_and_dynamic():
    a, b = np.array([-3.0, 0, 0, 0]), np.array([3.0, 0, 0, 0])
    tau, _ = ob.steer(a, b)
    obs = np.array([[0.0, 0.5, 0.0, 0.0]])
    assert ob.edge_collide(a, b, tau[0], 0.0, obs, 0) == 1
    assert ob.edge_collide(a, b, tau[0], 0.0, np.array([[0.0, 3.0, 0, 0]]), 0) == 0
    # obstacle that is on the path at t=0 but has moved away by the time the robot gets there
    mov = np.array([[0.0, 0.0, 0.0, 5.0]])
    assert ob.edge_collide(a, b, tau[0], 0.0, mov, 0) == 1
    assert ob.edge_collide(a, b, tau[0], 0.0, mov, 1) == 0
    # ... and one that arrives exactly when the robot does
    h = 0.5
    arr = np.array([[0.0, -1.0 * tau[0] * h, 0.0, 1.0]])
    assert ob.edge_collide(a, b, tau[0], 0.0, arr, 1) == 1
    assert ob.edge_collide(a, b, tau[0], 100.0, arr, 1) == 0


@pytest.mark.parametrize("dynamic,p_direct", [(0, 0.0), (1, 0.1)])
def test_kdtree_equals_brute_force_and_invariants(dynamic, p_direct):
    trees = []
    for kd in (1, 0):
        P = ob.OraclePlanner(capacity=6000, p_direct=p_direct, use_kdtree=kd)
        P.set_obstacles(scenes.dynamic_obstacles() if dynamic else scenes.static_obstacles(), dynamic)
        P.set_start(scenes.START)
        P.grow(scenes.GOAL, 4000)
        assert P.check_invariants() == 0
        trees.append((P.tree(), P.stats()))
    for x, y in zip(trees[0][0], trees[1][0]):
        assert np.array_equal(x, y)
    assert trees[0][1] == trees[1][1]


def test_stored_edges_are_collision_free_static():
    P = ob.OraclePlanner(capacity=3000)
    obs = scenes.static_obstacles()
    P.set_obstacles(obs, 0)
    P.set_start(scenes.START)
    P.grow(None, 2000)
    st, par, ctc, tim = P.tree()
    for i in range(1, len(par)):
        tau, _ = ob.steer(st[par[i]], st[i])
        assert ob.edge_collide(st[par[i]], st[i], tau[0], tim[par[i]], obs, 0) == 0


def test_cost_range_and_improvement_like_readme():
    """README cost_plot.png: on an 18 x 12 field costs sit in the 11-16 band and fall as the tree grows.  The
    README scene is unknown; here the obstacle-free optimum for START->GOAL is the rest-to-rest closed form
    (10.81, the bottom of that band) and the 9-robot slalom scene costs more, never less, and improves with n."""
    d = np.linalg.norm(scenes.GOAL[:2] - scenes.START[:2])
    tau = (36 * 0.5 * d * d) ** 0.25
    best = tau + 36 * 0.5 * d * d / 3 / tau**3
    P = ob.OraclePlanner(capacity=512, p_direct=0.1)
    P.set_obstacles(np.zeros((0, 4)), 0)
    P.set_start(scenes.START)
    ok, gi = P.grow(scenes.GOAL, 300)
    assert ok and abs(P.result(ok, gi)["cost"] - best) < 1e-9 * best
    costs = []
    for n_iter in (200, 2000):
        cs = []
        for seed in range(6):
            P = ob.OraclePlanner(capacity=n_iter + 2, p_direct=0.1, seed=seed)
            P.set_obstacles(scenes.static_obstacles(), 0)
            P.set_start(scenes.START)
            ok, gi = P.grow(scenes.GOAL, n_iter)
            assert ok
            cs.append(P.result(ok, gi)["cost"])
        costs.append(np.array(cs))
    assert np.all(costs[1] <= costs[0]) and costs[1].mean() < costs[0].mean()
    assert best < costs[1].min() and costs[0].max() < 60.0


def test_golden_vectors():
    g = np.load(os.path.join(GOLD, "steer.npz"))
    t, c = ob.steer(g["a"], g["b"], float(g["r"]))
    assert np.array_equal(t, g["tau"]) and np.array_equal(c, g["cost"])
    g = np.load(os.path.join(GOLD, "tree_c2_small.npz"))
    P = ob.OraclePlanner(capacity=int(g["capacity"]), p_direct=float(g["p_direct"]))
    P.set_obstacles(g["obs"], 1)
    P.set_start(g["start"])
    P.grow(g["goal"], int(g["n_iter"]))
    st, par, ctc, tim = P.tree()
    assert np.array_equal(par, g["parent"]) and np.array_equal(st, g["states"])
    assert np.array_equal(ctc, g["ctc"]) and np.array_equal(tim, g["time"])
    g = np.load(os.path.join(GOLD, "batch_small.npz"))
    out = ob.plan_batch(ob.default_config(p_direct=0.1, seed=5), g["starts"], g["goals"], g["obs"], 1,
                        int(g["n_iter"]), query_base=int(g["query_base"]), nthreads=2)
    assert np.array_equal(out.view(np.uint8), g["results"])
