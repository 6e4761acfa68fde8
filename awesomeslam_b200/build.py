"""Builds libaslam_b200.so (CUDA, sm_100a) in-tree and, for the tests, the CPU oracle.

Everything is compiled with explicit nvcc / g++ command lines; the products are
  awesomeslam_b200/lib/libaslam_b200.so   the C-ABI library (include/aslam_c.h)
  awesomeslam_b200/lib/aslam_replay       the ROS-free C++ replay harness (include/aslam_filters.hpp)
  oracle/liboracle.so                     TEST INFRASTRUCTURE: CPU restatement of the reference
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "awesomeslam_b200", "csrc")
LIBDIR = os.path.join(ROOT, "awesomeslam_b200", "lib")
LIB = os.path.join(LIBDIR, "libaslam_b200.so")
ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-fmad=false",  # no implicit FMA contraction: hot loops call fma() explicitly
    "-Xcompiler", "-fPIC,-O2,-pthread", "-cudart", "static",
]


def _sources():
    out = []
    for name in sorted(os.listdir(CSRC)):
        if name.endswith(".cu") or name.endswith(".cpp"):
            out.append(os.path.join(CSRC, name))
    return out


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _all_deps():
    deps = []
    for d in (CSRC, os.path.join(ROOT, "include")):
        for name in os.listdir(d):
            deps.append(os.path.join(d, name))
    return deps


def build_library(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    headers = [p for p in _all_deps() if p.endswith((".h", ".hpp", ".cuh"))]
    objs = []
    procs = []
    for src in _sources():
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", src, "-o", obj]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(out)
        if p.returncode != 0:
            failed = True
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or not os.path.exists(LIB):
        cmd = ["nvcc", "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-pthread", "-o", LIB] + objs + ["-ldl"]
        subprocess.check_call(cmd)
    return LIB


HOST_LIB = os.path.join(LIBDIR, "libaslam_host.so")


def build_host_lib(force=False):
    """csrc/replay.cpp alone (replay generator + quat2euler): host-only helpers without CUDA."""
    os.makedirs(LIBDIR, exist_ok=True)
    src = os.path.join(CSRC, "replay.cpp")
    if force or _stale(HOST_LIB, [src, os.path.join(ROOT, "include", "aslam_c.h")]):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", src, "-o", HOST_LIB])
    return HOST_LIB


HARNESS = os.path.join(LIBDIR, "aslam_replay")


def build_harness(force=False):
    """The ROS-free C++ replay harness (awesomeslam_b200/harness/replay_main.cpp) over include/aslam_filters.hpp."""
    src = os.path.join(ROOT, "awesomeslam_b200", "harness", "replay_main.cpp")
    deps = [src, os.path.join(ROOT, "include", "aslam_filters.hpp"), os.path.join(ROOT, "include", "aslam_c.h"), LIB]
    if force or _stale(HARNESS, deps):
        subprocess.check_call(["g++", "-O2", "-std=c++14", "-I", os.path.join(ROOT, "include"), src, "-L", LIBDIR,
                               "-laslam_b200", "-Wl,-rpath,$ORIGIN", "-pthread", "-ldl", "-lrt", "-o", HARNESS])
    return HARNESS


def build_oracle(force=False):
    odir = os.path.join(ROOT, "oracle")
    srcs = [os.path.join(odir, "oracle_c.cpp")]
    deps = [os.path.join(odir, n) for n in os.listdir(odir) if n.endswith((".cpp", ".hpp"))]
    if force or _stale(ORACLE_LIB, deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
                               "-o", ORACLE_LIB] + srcs)
    return ORACLE_LIB


REF_SRC = "/root/reference/awesome_slam"
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def build_ref(force=False):
    """TEST INFRASTRUCTURE: compiles the reference's own ekf.cpp / ukf.cpp / sensor_landmark.cpp, unmodified and from
    where they lie under /root/reference, against the Eigen / ROS stand-ins of oracle/ref_shim into oracle/_ref/
    (git-ignored, travels to the GPU box). Where /root/reference is absent (the GPU box) the prebuilt files are
    used as they are. Returns the directory, or None when neither sources nor prebuilt libraries exist."""
    if os.path.isdir(REF_SRC):
        cmd = ["make", "-C", os.path.join(ROOT, "oracle", "ref_shim"), "REF=" + REF_SRC] + (["-B"] if force else [])
        subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
        return REF_DIR
    return REF_DIR if os.path.exists(os.path.join(REF_DIR, "libref_ekf.so")) else None


if __name__ == "__main__":
    v = "-v" in sys.argv
    print(build_library(force="-f" in sys.argv, verbose=v))
    print(build_host_lib(force="-f" in sys.argv))
    print(build_harness(force="-f" in sys.argv))
    print(build_oracle(force="-f" in sys.argv))
    print(build_ref(force="-f" in sys.argv))
