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
