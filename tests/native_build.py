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


REF_PKG = "/root/reference/src/robosoccer_motion_planning"


def build_reference_node():
    """The reference's OWN node — src/robosoccer_motion_planning_node.cpp + the package library's four sources, all
    compiled UNMODIFIED where they lie under /root/reference (nothing is copied) — against include/*.hpp of this repo,
    with tests/native/ros_stub/ standing in for roscpp / tf / the message headers.  Same structure as the reference's
    CMakeLists.txt:131-165: a shared library (its headers define non-inline functions, so it only links as one) and the
    node executable.  g++ -std=c++17 -fconcepts (the sources use `auto` parameters, CMakeLists.txt:5 + GCC's extension);
    the one force-included shim is the std::__cxx11::string typedef trajectory1d.cpp:196 needs (oracle/Makefile).
    Returns the executable, or None when neither /root/reference nor a prebuilt binary is there (the GPU box runs the
    binary built in the build container)."""
    from robosoccer_motion_planning_b200 import capi
    capi.build_library()
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "reference_node")
    so = os.path.join(OUT, "librobosoccer_motion_planning.so")
    if not os.path.isdir(os.path.join(REF_PKG, "src")):
        return exe if os.path.exists(exe) and os.path.exists(so) else None
    stub = os.path.join(ROOT, "tests", "native", "ros_stub")
    inc = os.path.join(ROOT, "include")
    deps = [os.path.join(inc, f) for f in ("rrtstar.hpp", "integrator2drrt.hpp", "rrtvisual.hpp", "krrt.h")] + [capi.LIB_PATH]
    for d, _, fs in os.walk(stub):
        deps += [os.path.join(d, f) for f in fs]
    if os.path.exists(exe) and os.path.exists(so) and all(os.path.getmtime(exe) >= os.path.getmtime(d) for d in deps):
        return exe
    flags = ["-std=c++17", "-fconcepts", "-O2", "-w", "-include", os.path.join(ROOT, "oracle", "ref_wrap", "cxx11_string_shim.h"),
             "-I" + stub, "-I" + inc, "-I" + os.path.join(REF_PKG, "include")]
    lib_src = [os.path.join(REF_PKG, "src", f) for f in ("robotsubscriber.cpp", "goalsubscriber.cpp", "trajectorypublisher.cpp",
                                                          "trajectory1d.cpp")]
    subprocess.check_call(["g++"] + flags + ["-fPIC", "-shared", "-o", so] + lib_src + [os.path.join(stub, "ros_stub_runtime.cpp")])
    subprocess.check_call(["g++"] + flags + ["-o", exe, os.path.join(REF_PKG, "src", "robosoccer_motion_planning_node.cpp"),
                           "-L" + OUT, "-lrobosoccer_motion_planning", "-L" + os.path.dirname(capi.LIB_PATH), "-lkrrt", "-pthread",
                           "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/../../../robosoccer_motion_planning_b200/lib"])
    return exe


def build_benchmark():
    """robosoccer_motion_planning_b200/tools/rrtstar_benchmark.cpp (product tool: links libkrrt only)."""
    from robosoccer_motion_planning_b200 import capi
    capi.build_library()
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "rrtstar_benchmark")
    src = os.path.join(ROOT, "robosoccer_motion_planning_b200", "tools", "rrtstar_benchmark.cpp")
    deps = [src, os.path.join(ROOT, "include", "rrtstar.hpp"), os.path.join(ROOT, "include", "integrator2drrt.hpp"),
            os.path.join(ROOT, "include", "krrt.h"), capi.LIB_PATH]
    if os.path.exists(exe) and all(os.path.getmtime(exe) >= os.path.getmtime(d) for d in deps):
        return exe
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-o", exe, src,
                           "-L" + os.path.dirname(capi.LIB_PATH), "-lkrrt", "-Wl,-rpath," + os.path.dirname(capi.LIB_PATH),
                           "-Wl,-rpath,$ORIGIN/../../../robosoccer_motion_planning_b200/lib"])
    return exe


def build_cmake_package(build_dir=None):
    """The `rrtstar` package through its own CMakeLists.txt (plain CMake, no catkin): libkrrt.so, libkrrt_mgpu.so and the two
    benchmark programs.  Returns the build directory."""
    import shutil
    build_dir = build_dir or os.path.join(ROOT, "build", "cmake")
    cmake = shutil.which("cmake")
    if cmake is None:
        return None
    args = [cmake, "-S", ROOT, "-B", build_dir]
    if shutil.which("ninja"):
        args += ["-G", "Ninja"]
    try:
        import nvidia.nccl as n
        args.append("-DNCCL_HOME=" + list(n.__path__)[0])
    except ImportError:
        pass
    if not os.path.exists(os.path.join(build_dir, "CMakeCache.txt")):
        subprocess.check_call(args, stdout=subprocess.DEVNULL)
    subprocess.check_call([cmake, "--build", build_dir, "-j", "8"], stdout=subprocess.DEVNULL)
    return build_dir
