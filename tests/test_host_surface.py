"""The C++17 component surface (include/rrtstar.hpp, include/integrator2drrt.hpp): a caller written like the
reference's node (tests/native/host_surface_check.cpp) must compile with plain g++ -std=c++17, fail loudly without a
GPU, and on a B200 grow exactly the oracle's tree through one grow(xg) call per iteration."""
import subprocess

import pytest

import native_build


def test_surface_compiles_and_has_no_cpu_fallback():
    exe = native_build.build_host_surface_check()
    from robosoccer_motion_planning_b200 import capi
    if capi.lib().krrt_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "static", "20"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3, r.stdout + r.stderr          # krrt::Error(KRRT_ERR_NO_DEVICE)
    assert "no CPU fallback" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("env,iters", [("static", 100), ("dynamic", 100), ("dynamic", 400)])
def test_node_style_caller_matches_oracle(gpu_required, env, iters):
    exe = native_build.build_host_surface_check()
    r = subprocess.run([exe, env, str(iters)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 check failure(s)" in r.stdout


def test_service_compiles_and_has_no_cpu_fallback():
    exe = native_build.build_service_check()
    from robosoccer_motion_planning_b200 import capi
    if capi.lib().krrt_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3, r.stdout + r.stderr
    assert "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_trajectory_service_matches_checkers(gpu_required):
    """compute_trajectory (trajectory_generator_node.cpp:6-53) for "angular_trajectory" and the 2-D integrator model."""
    exe = native_build.build_service_check()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 check failure(s)" in r.stdout


def test_custom_benchmark_builds_and_fails_loudly_without_gpu():
    exe = native_build.build_custom_benchmark()
    from robosoccer_motion_planning_b200 import capi
    if capi.lib().krrt_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "1"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_custom_benchmark_runs(gpu_required):
    """The reference's `rrtstar_custom_benchmark` (README.md:39-42) over the component surface: 4 sampling settings x 7
    node sizes, min / mean / max of time and cost; with direct sampling p >= 0.1 and 200 nodes most trials find the goal."""
    exe = native_build.build_custom_benchmark()
    r = subprocess.run([exe, "8"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    rows = [ln.split() for ln in r.stdout.splitlines() if ln.startswith("  ")]
    assert len(rows) == 28
    last = [w for w in rows if w[0] == "0.50" and w[1] == "200"][0]
    assert int(last[5].split("/")[0]) >= 6 and float(last[7]) > 0.0
    assert float(last[8]) > float(last[6])                     # trials differ (one sample stream per trial)


def test_benchmark_builds_and_fails_loudly_without_gpu():
    exe = native_build.build_benchmark()
    from robosoccer_motion_planning_b200 import capi
    if capi.lib().krrt_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_benchmark_runs(gpu_required):
    """The reference's `rrtstar_benchmark` (README.md:34-37): google-benchmark conventions, config 1 of BASELINE.json."""
    import json
    exe = native_build.build_benchmark()
    r = subprocess.run([exe, "--benchmark_min_time=0.05", "--benchmark_format=json"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    b = {x["name"]: x for x in json.loads(r.stdout)["benchmarks"]}
    assert len(b) == 9 and b["BM_RRTStarInt2D/10000"]["solved"] == 1 and b["BM_RRTStarInt2D/10000"]["nodes"] > 5000
    assert all(x["iterations"] >= 3 and x["real_time"] > 0 for x in b.values())
    # one plan of 75 iterations and the caller's 75 single calls grow the same tree
    assert b["BM_RRTStarInt2D/75"]["nodes"] == b["BM_RRTStarInt2D_PerCall/75"]["nodes"]
    assert b["BM_RRTStarInt2D/75"]["cost"] == b["BM_RRTStarInt2D_PerCall/75"]["cost"]


def test_cmake_package_builds():
    """CMakeLists.txt + package.xml of the `rrtstar` package: plain CMake builds both libraries and both tools, and the
    library exports the C ABI."""
    import ctypes
    import os
    d = native_build.build_cmake_package()
    if d is None:
        pytest.skip("cmake not installed")
    for f in ("libkrrt.so", "libkrrt_mgpu.so", "rrtstar_benchmark", "rrtstar_custom_benchmark"):
        assert os.path.exists(os.path.join(d, f)), f
    L = ctypes.CDLL(os.path.join(d, "libkrrt.so"))
    from robosoccer_motion_planning_b200 import capi
    for name in capi.EXPORTS:
        assert hasattr(L, name), name
