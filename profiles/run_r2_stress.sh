#!/bin/bash
# Round 2: stability of the final state — the GPU suite three times over, the bench twice, 200 repeats of the C2 plan and of a
# logged 75-iteration cycle (every run compared with the first), under a watchdog.
set -u
mkdir -p gpurun_out
for k in 1 2 3; do timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -1; done > gpurun_out/stress_pytest.log 2>&1; cat gpurun_out/stress_pytest.log
for k in 1 2; do timeout 600 python bench.py --skip-c3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['parity_vs_oracle'], d['batched']['queries_per_sec'], d['batched']['parity_vs_oracle'], d['clocks'])"; done
timeout 600 python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests')
import numpy as np
from robosoccer_motion_planning_b200 import capi, scenes
G = capi.Planner(capacity=20002, p_direct=0.1)
G.set_obstacles(scenes.dynamic_obstacles(), 1)
ref = None; ms = []
for k in range(200):
    G.set_start(scenes.START)
    r = G.grow(scenes.GOAL, 20000)
    ms.append(G.last_timing()['total_ms'])
    t = G.tree() + (np.array(list(G.stats().values())),)
    if ref is None: ref = t
    else: assert all(np.array_equal(a, b) for a, b in zip(t, ref)), k
print('200 C2 plans identical; ms min/median/max', min(ms), sorted(ms)[100], max(ms))
ref = None
for k in range(300):
    G.set_start(scenes.START)
    rec, s, gi = G.grow_logged(scenes.GOAL, 75)
    if ref is None: ref = rec.copy()
    else: assert rec.tobytes() == ref.tobytes(), k
print('300 logged cycles identical')
PY
