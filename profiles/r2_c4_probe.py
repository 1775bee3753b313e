#!/usr/bin/env python
"""C4 probe (4096 queries x 200 iterations, 9 moving obstacles, p_direct 0.1, inputs resident in HBM): device time of the
batched planner per precision, through libkrrt_mgpu on one GPU; a 256-query sample is compared with the matching CPU checker."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from robosoccer_motion_planning_b200 import capi, scenes, sweep  # noqa: E402

q, iters = int(os.environ.get("Q", 4096)), int(os.environ.get("ITERS", 200))
bs, bg, bo = scenes.batch_queries(q, 9, seed=0)
for name, prec in (("f64", capi.KRRT_F64), ("f32", capi.KRRT_F32)):
    S = sweep.Sweep.single_process(1, p_direct=0.1, precision=prec)
    S.load(bs, bg, bo)
    for _ in range(3):
        S.run(1, iters)
    ms = [S.run(1, iters) for _ in range(7)]
    out = S.results()
    st = S.stats()
    if prec == capi.KRRT_F32:
        import oracle_f32_binding as of
        ref = of.plan_batch(of.default_config(p_direct=0.1), bs[:256], bg[:256], bo[:256], 1, iters, nthreads=8)
    else:
        import oracle_binding as ob
        ref = ob.plan_batch(ob.default_config(p_direct=0.1), bs[:256], bg[:256], bo[:256], 1, iters, nthreads=8)
    same = np.array_equal(out[:256].view(np.uint8), ref.view(np.uint8))
    print(f"{name}: best {min(ms):.3f} ms  median {sorted(ms)[len(ms) // 2]:.3f} ms  {q / min(ms) / 1e3:.3f} M queries/s  "
          f"solved {out['solved'].mean():.4f}  mean cost {out['cost'][out['solved'] != 0].mean():.4f}  "
          f"parity(256) {'bit-exact' if same else 'MISMATCH'}  steer {st['steer_calls']} collide {st['collide_calls']}")
    S.close()
