"""Synthetic soccer-field scenes for the configs of BASELINE.json.

Field 18 x 12 m (gazebo_visual/.../global_config.yaml:17-18); robots spawn at the poses of
robot_up.sh:11-14; input weight r = 0.5 and collision radius 1.35 from
robosoccer_motion_planning/config/config.yaml:1,5; direct-goal sampling p = 0.1 (config.yaml:7).
"""
import numpy as np

FIELD_LO = np.array([-9.0, -6.0, -2.0, -2.0])
FIELD_HI = np.array([9.0, 6.0, 2.0, 2.0])
START = np.array([-8.5, 0.0, 0.0, 0.0])
GOAL = np.array([7.0, 0.0, 0.0, 0.0])
R_COLLISION = 1.35
SPAWN_XY = np.array([(-2, 1), (-2, -1), (-5, 2), (-5, -2), (8.5, 0), (2, 1), (2, -1), (5, 2), (5, -2)], dtype=np.float64)


def static_obstacles():
    """C1: the 9 other robots standing at their spawn poses."""
    obs = np.zeros((9, 4))
    obs[:, :2] = SPAWN_XY
    return obs


def dynamic_obstacles(seed=1):
    """C2: same positions, velocities ~ U(-1,1)^2 m/s."""
    rng = np.random.default_rng(seed)
    obs = static_obstacles()
    obs[:, 2:] = rng.uniform(-1.0, 1.0, (9, 2))
    return obs


def batch_queries(q, m=9, seed=0, first=0):
    """C4/C5: q independent queries.  start, goal ~ U(field) with zero velocity; m moving discs ~ U(field)
    kept at least R from start and goal, v ~ U(-1,1)^2.  Query i depends only on (seed, first + i), so a
    shard of a larger sweep generates exactly its own slice."""
    starts = np.zeros((q, 4))
    goals = np.zeros((q, 4))
    obs = np.zeros((q, m, 4))
    for i in range(q):
        rng = np.random.default_rng([seed, first + i])
        starts[i, :2] = rng.uniform(FIELD_LO[:2], FIELD_HI[:2])
        goals[i, :2] = rng.uniform(FIELD_LO[:2], FIELD_HI[:2])
        for j in range(m):
            while True:
                p = rng.uniform(FIELD_LO[:2], FIELD_HI[:2])
                if (np.hypot(*(p - starts[i, :2])) >= 1.5 * R_COLLISION
                        and np.hypot(*(p - goals[i, :2])) >= 1.5 * R_COLLISION):
                    break
            obs[i, j, :2] = p
            obs[i, j, 2:] = rng.uniform(-1.0, 1.0, 2)
    return starts, goals, obs
