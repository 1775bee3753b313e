"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/krrt.h declares,
and FAILS LOUDLY without a GPU (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

from robosoccer_motion_planning_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    capi.build_library()
    return capi.lib()


def header_functions():
    src = open(os.path.join(ROOT, "include", "krrt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(krrt_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(built, n), f"libkrrt.so does not export {n}"
    assert sorted(capi.EXPORTS) == names


def test_struct_layouts_match_header(built):
    assert ctypes.sizeof(capi.KrrtConfig) == 4 * 4 + 2 * 8 + 8 * 8 + 4 * 8 + 8 + 4 + 4 + 4 + 4
    assert capi.RESULT_DTYPE.itemsize == 16 + 16 + 32 * 5 * 8
    c = capi.default_config()
    assert (c.r, c.R, c.S) == (0.5, 1.35, 10)           # config/config.yaml:1,5
    assert list(c.lo) == [-9, -6, -2, -2] and list(c.hi) == [9, 6, 2, 2]


def test_no_cpu_fallback(built):
    """Without a CUDA device creation must fail with KRRT_ERR_NO_DEVICE — never a silent host path."""
    if built.krrt_device_count() > 0:
        pytest.skip("a GPU is visible; the failure path is for GPU-less hosts")
    with pytest.raises(capi.KrrtError) as e:
        capi.Planner()
    assert e.value.code == capi.KRRT_ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_invalid_config_rejected(built):
    for kw in (dict(capacity=1), dict(batch=2000), dict(near_cap=7), dict(r=0.0), dict(precision=2)):
        with pytest.raises(capi.KrrtError) as e:
            capi.Planner(**kw)
        assert e.value.code in (capi.KRRT_ERR_INVALID, capi.KRRT_ERR_UNSUPPORTED)


def test_product_does_not_touch_the_oracle():
    """Nothing under the package or include/ may reference oracle/ (the oracle is test infrastructure)."""
    for top in ("robosoccer_motion_planning_b200", "include"):
        for base, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", ".hpp", "Makefile")):
                    txt = open(os.path.join(base, f), errors="ignore").read()
                    for word in ("krrt_oracle", "yaw_oracle", "ingest_oracle", "oracle_binding", "yaw_binding",
                                 "ingest_binding", "oracle/"):
                        assert word not in txt, (os.path.join(base, f), word)
    out = os.popen(f"ldd {capi.LIB_PATH}").read()
    assert "oracle" not in out


def test_scene_generators_are_shard_invariant():
    from robosoccer_motion_planning_b200 import scenes
    s, g, o = scenes.batch_queries(10, 9, seed=3)
    s2, g2, o2 = scenes.batch_queries(4, 9, seed=3, first=6)
    assert np.array_equal(s[6:], s2) and np.array_equal(g[6:], g2) and np.array_equal(o[6:], o2)
    d = np.linalg.norm(o[:, :, :2] - s[:, None, :2], axis=2)
    assert d.min() >= 1.5 * scenes.R_COLLISION
