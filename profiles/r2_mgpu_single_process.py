#!/usr/bin/env python
"""libkrrt_mgpu driven by ONE process (krrt_mgpu_create: a host thread + context + stream per GPU, ncclCommInitAll): a sweep over
all GPUs of the box against the same sweep on one GPU (records byte-equal) and the per-sweep device time."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from robosoccer_motion_planning_b200 import capi, scenes, sweep  # noqa: E402

n = capi.lib().krrt_device_count()
q = int(os.environ.get("Q", 16384))
s, g, o = scenes.batch_queries(4096, 9, seed=0)
sel = np.arange(q) % 4096
s, g, o = (np.ascontiguousarray(x[sel]) for x in (s, g, o))
ref = None
for k in sorted({1, n}):
    S = sweep.Sweep.single_process(k, p_direct=0.1)
    out = S.sweep(s, g, o, 1, 200)
    S.load(s, g, o)
    S.run(1, 200)
    ms = min(S.run(1, 200) for _ in range(3))
    h = hashlib.sha256(out.tobytes()).hexdigest()[:16]
    print(f"{k} GPU(s) in one process: {q} queries, {ms:.3f} ms per sweep, {q / ms / 1e3:.2f} M queries/s, sha256 {h}, info {S.info()}, overlap {S.overlap()}")
    if ref is None:
        ref = out
    else:
        print("records equal to the 1-GPU sweep:", bool(np.array_equal(ref.view(np.uint8), out.view(np.uint8))))
    S.close()
