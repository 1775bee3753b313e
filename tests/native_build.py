"""Builds the native (C++) test programs under tests/native/_build (test infrastructure: they link the oracle)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "native", "_build")


def build_host_surface_check():
    from robosoccer_motion_planning_b200 import capi
    import oracle_binding as ob
    capi.build_library()
    ob.build_oracle()
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "host_surface_check")
    src = os.path.join(ROOT, "tests", "native", "host_surface_check.cpp")
    deps = [src, os.path.join(ROOT, "include", "rrtstar.hpp"), os.path.join(ROOT, "include", "integrator2drrt.hpp"),
            os.path.join(ROOT, "include", "krrt.h"), capi.LIB_PATH]
    if os.path.exists(exe) and all(os.path.getmtime(exe) >= os.path.getmtime(d) for d in deps):
        return exe
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-o", exe, src,
           "-L" + os.path.dirname(capi.LIB_PATH), "-lkrrt", "-L" + os.path.join(ROOT, "oracle", "_build"), "-lkrrt_oracle",
           "-Wl,-rpath," + os.path.dirname(capi.LIB_PATH), "-Wl,-rpath," + os.path.join(ROOT, "oracle", "_build"),
           "-Wl,-rpath,$ORIGIN/../../../robosoccer_motion_planning_b200/lib", "-Wl,-rpath,$ORIGIN/../../../oracle/_build"]
    subprocess.check_call(cmd)
    return exe


def build_service_check():
    """tests/native/service_check.cpp: include/trajectory_service.hpp against the CPU checkers."""
    from robosoccer_motion_planning_b200 import capi
    import oracle_binding as ob
    capi.build_library()
    ob.build_oracle(force=not os.path.exists(os.path.join(ROOT, "oracle", "_build", "libyaw_oracle.so")))
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "service_check")
    src = os.path.join(ROOT, "tests", "native", "service_check.cpp")
    odir = os.path.join(ROOT, "oracle", "_build")
    deps = [src, os.path.join(ROOT, "include", "trajectory_service.hpp"), os.path.join(ROOT, "include", "rrtstar.hpp"),
            os.path.join(ROOT, "include", "krrt.h"), capi.LIB_PATH, os.path.join(odir, "libkrrt_oracle.so"),
            os.path.join(odir, "libyaw_oracle.so")]
    if os.path.exists(exe) and all(os.path.getmtime(exe) >= os.path.getmtime(d) for d in deps):
        return exe
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-o", exe, src,
           "-L" + os.path.dirname(capi.LIB_PATH), "-lkrrt", "-L" + odir, "-lkrrt_oracle", "-lyaw_oracle",
           "-Wl,-rpath," + os.path.dirname(capi.LIB_PATH), "-Wl,-rpath," + odir,
           "-Wl,-rpath,$ORIGIN/../../../robosoccer_motion_planning_b200/lib", "-Wl,-rpath,$ORIGIN/../../../oracle/_build"]
    subprocess.check_call(cmd)
    return exe


def build_custom_benchmark():
    """robosoccer_motion_planning_b200/tools/rrtstar_custom_benchmark.cpp (product tool: links libkrrt only)."""
    from robosoccer_motion_planning_b200 import capi
    capi.build_library()
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "rrtstar_custom_benchmark")
    src = os.path.join(ROOT, "robosoccer_motion_planning_b200", "tools", "rrtstar_custom_benchmark.cpp")
    deps = [src, os.path.join(ROOT, "include", "rrtstar.hpp"), os.path.join(ROOT, "include", "integrator2drrt.hpp"),
            os.path.join(ROOT, "include", "krrt.h"), capi.LIB_PATH]
    if os.path.exists(exe) and all(os.path.getmtime(exe) >= os.path.getmtime(d) for d in deps):
        return exe
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-o", exe, src,
                           "-L" + os.path.dirname(capi.LIB_PATH), "-lkrrt", "-Wl,-rpath," + os.path.dirname(capi.LIB_PATH),
                           "-Wl,-rpath,$ORIGIN/../../../robosoccer_motion_planning_b200/lib"])
    return exe
