"""Worker of tests/test_switch_matrix.py (run as a subprocess so that the CHECKED build of the library is the one loaded):
every setting of the seven prefix-relaxation switches of k_grow on two sample streams, trees and counters against the oracle."""
import itertools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from robosoccer_motion_planning_b200 import capi  # noqa: E402

capi.LIB_PATH = os.path.join(capi.PKG_DIR, "lib", "libkrrt_check.so")
capi.build_library = lambda *a, **k: capi.LIB_PATH
import oracle_binding as ob  # noqa: E402
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

SWITCHES = ("KRRT_PEND", "KRRT_RAW_FORWARD", "KRRT_PROP_FORWARD", "KRRT_JOIN_TEST", "KRRT_INTRA_FIX", "KRRT_TAIL_SEQ", "KRRT_FLOW")
n_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
stride = int(sys.argv[2]) if len(sys.argv) > 2 else 1
bad = 0
for query_id in (0, 5):
    O = ob.OraclePlanner(capacity=n_iter + 2, p_direct=0.1, query_id=query_id)
    O.set_obstacles(scenes.dynamic_obstacles(), 1)
    O.set_start(scenes.START)
    ro = O.grow(scenes.GOAL, n_iter)
    ot, ost = O.tree(), O.stats()
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1, query_id=query_id)
    G.set_obstacles(scenes.dynamic_obstacles(), 1)
    for k, bits in enumerate(itertools.product((0, 1), repeat=len(SWITCHES))):
        if k % stride:
            continue
        for name, b in zip(SWITCHES, bits):
            os.environ[name] = str(b)
        G.set_start(scenes.START)
        rg = G.grow(scenes.GOAL, n_iter)
        same = rg == ro and all(np.array_equal(a, b) for a, b in zip(G.tree(), ot)) and G.stats() == ost
        if not same:
            bad += 1
            print("MISMATCH query", query_id, dict(zip(SWITCHES, bits)))
    G.close()
print("switch matrix:", 2 * (2 ** len(SWITCHES) // stride), "plans of", n_iter, "iterations on the checked build,", bad, "mismatch(es)")
sys.exit(1 if bad else 0)
