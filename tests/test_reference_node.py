"""The reference's own planner node, unmodified, on this repo's component set.

/root/reference/src/robosoccer_motion_planning/src/robosoccer_motion_planning_node.cpp (with include/robosoccer_visual.hpp
and the package library's sources) is compiled where it lies against include/rrtstar.hpp, integrator2drrt.hpp and
rrtvisual.hpp; roscpp / tf / message headers are stand-ins under tests/native/ros_stub (test infrastructure).  On a B200
the node plans one cycle on the C1 / C2 scene through its own `for(i<target_size) rrt.grow(xg)` loop (…node.cpp:176-185),
re-solves the path at dt = 0.1 (:232-249) and publishes it; the published nav_msgs/Path must equal, bit for bit, what the
CPU oracle gives for the same calls."""
import os
import subprocess

import numpy as np
import pytest

import native_build


def _run(exe, env_extra, timeout=300):
    env = dict(os.environ)
    env.update(env_extra)
    return subprocess.run([exe], capture_output=True, text=True, timeout=timeout, env=env)


def test_reference_node_compiles_unmodified_and_has_no_cpu_fallback():
    exe = native_build.build_reference_node()
    if exe is None:
        pytest.skip("/root/reference is not mounted and no prebuilt node is present")
    from robosoccer_motion_planning_b200 import capi
    if capi.lib().krrt_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = _run(exe, {})
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr          # krrt::Error(KRRT_ERR_NO_DEVICE) out of rrt.setStart: nothing plans on the host


def _oracle_cycle(dynamic, target_size, direct, p, f32=False):
    """What …node.cpp:118-141,167-196,232-249 does in one cycle, on the oracle (its float build for a binary32 run)."""
    import oracle_binding as ob
    from robosoccer_motion_planning_b200 import scenes
    if f32:
        import oracle_f32_binding as of
        ob = type("F32", (), {"OraclePlanner": staticmethod(lambda **kw: of.OraclePlannerF32(**kw)), "solve": staticmethod(of.solve)})
    vel = np.array([(0.5, -0.25), (-0.5, 0.75), (0.25, 0.5), (-0.75, -0.5), (-1, 0.25), (0.5, 0.5), (-0.25, -0.75), (0.75, 0.25),
                    (-0.5, -0.25)])
    obs = scenes.static_obstacles()
    obs[:, 2:] = vel if dynamic else 0.0
    O = ob.OraclePlanner(capacity=1 << 16, p_direct=p if direct else 0.0)
    O.set_obstacles(obs, 1 if dynamic else 0)
    O.set_start(scenes.START)
    goal = scenes.GOAL
    forced = False
    solved, gi = False, -1
    for i in range(target_size):
        row = goal.reshape(1, 4) if forced else None
        solved, gi = O.grow(goal, 1, samples=row)
        forced = direct and (i == target_size - 2 or i == target_size // 2)
    if not solved:
        return None, O
    states = O.path(gi)[0]
    rows = []
    t0 = 0.0
    for a, b in zip(states[:-1], states[1:]):
        seg = ob.solve(a, b, 0.5, 0.1, 0.0)
        seg[:, 4] += t0
        rows.append(seg)
        t0 = seg[-1, 4]
    return np.concatenate(rows), O


@pytest.mark.gpu
@pytest.mark.parametrize("environment,target_size,direct,precision",
                         [("static", 100, False, "f64"), ("dynamic", 75, True, "f64"), ("dynamic", 200, True, "f64"),
                          ("dynamic", 75, True, "f32")])
def test_reference_node_publishes_the_oracles_path(gpu_required, tmp_path, environment, target_size, direct, precision):
    exe = native_build.build_reference_node()
    if exe is None:
        pytest.skip("no reference node binary (built where /root/reference is mounted; it travels to the GPU box)")
    base = str(tmp_path / "path")
    vis = str(tmp_path / "vis.csv")
    env = {"KRRT_STUB_CYCLES": "1", "KRRT_STUB_GOAL": "7,0", "KRRT_STUB_PATH_CSV": base, "KRRT_RRTVISUAL_CSV": vis,
           "ROSPARAM_environment": environment, "ROSPARAM_target_size": str(target_size), "ROSPARAM_rrtvis_rate": "1e9",
           "ROSPARAM_direct_sampling": "true" if direct else "false", "ROSPARAM_direct_sampling_prob": "0.1",
           "KRRT_PRECISION": precision}
    r = _run(exe, env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "1 cycle(s) done" in r.stdout
    want, O = _oracle_cycle(environment == "dynamic", target_size, direct, 0.1, f32=precision == "f32")
    assert ("solution" in r.stderr) and (("not found" not in r.stderr) == (want is not None)), r.stderr
    # the recorder saw the tree: one frame, a DELETEALL in front, the tree's edges
    frame = np.genfromtxt(vis, delimiter=",", skip_header=1, usecols=(0, 1, 2))
    assert frame.shape[0] > O.size and int(frame[0, 2]) == 3
    if want is None:
        return
    pos = np.genfromtxt(base + ".robosoccer_trajectory_pos.csv", delimiter=",", skip_header=1)
    vel = np.genfromtxt(base + ".robosoccer_trajectory_vel.csv", delimiter=",", skip_header=1)
    assert pos.shape[0] == want.shape[0]
    assert np.array_equal(pos[:, 2:4], want[:, 0:2]) and np.array_equal(vel[:, 2:4], want[:, 2:4])
    # stamps: rrt_t0 + time since the root (…node.cpp:319-322)
    assert np.allclose(pos[:, 1] - pos[0, 1], want[:, 4] - want[0, 4], rtol=0, atol=1e-6)
    assert np.allclose(pos[-1, 2:4], [7.0, 0.0], rtol=0, atol=1e-9 if precision == "f64" else 1e-5)
