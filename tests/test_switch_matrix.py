"""Hardening of k_grow: the seven relaxations of the prefix rule are independent switches
(KRRT_PEND, KRRT_RAW_FORWARD, KRRT_PROP_FORWARD, KRRT_JOIN_TEST, KRRT_INTRA_FIX, KRRT_TAIL_SEQ, KRRT_FLOW); no combination may
be left that has never run.  All 2^7 settings x two sample streams are planned on the CHECKED build of the library
(-DKRRT_CHECK: every capacity the parallel commit relies on is asserted on the device and traps with a code) and compared, tree
and counters, with the oracle."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _check_lib():
    from robosoccer_motion_planning_b200 import capi
    subprocess.check_call(["make", "-C", capi.PKG_DIR, "-s", "check"])
    return os.path.join(capi.PKG_DIR, "lib", "libkrrt_check.so")


def test_checked_build_compiles_and_exports_the_abi():
    import ctypes
    from robosoccer_motion_planning_b200 import capi
    L = ctypes.CDLL(_check_lib())
    for name in capi.EXPORTS:
        assert hasattr(L, name), name


@pytest.mark.gpu
def test_every_switch_setting_on_the_checked_build(gpu_required):
    _check_lib()
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "switch_matrix_worker.py"), "8000"], capture_output=True,
                       text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "256 plans" in r.stdout and "0 mismatch(es)" in r.stdout


@pytest.mark.gpu
def test_large_tree_on_the_checked_build(gpu_required):
    """100k iterations (grid index, rebuilds, windows of 1024 samples, sub-rounds of up to 147 samples) with the device asserts on."""
    _check_lib()
    code = ("import os, sys; sys.path.insert(0, %r); from robosoccer_motion_planning_b200 import capi, scenes; import os.path as p;"
            "capi.LIB_PATH = p.join(capi.PKG_DIR, 'lib', 'libkrrt_check.so'); capi.build_library = lambda *a, **k: capi.LIB_PATH;"
            "G = capi.Planner(capacity=100002); G.set_obstacles(scenes.static_obstacles(), 0); G.set_start(scenes.START);"
            "G.grow(None, 100000); print('nodes', G.size, G.stats()['inserted'])") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "nodes 7" in r.stdout, r.stdout + r.stderr
