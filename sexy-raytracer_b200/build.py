"""In-tree build of the CUDA library (sm_100a only) and of the checkers under oracle/."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsrt_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              # parity contract: no FMA contraction anywhere but the explicit fmaf of the filtered box test;
              # IEEE division / sqrt and denormals are nvcc's defaults and must stay
              "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
              "-Xcompiler", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include")]


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_cuda(force=False, verbose=False):
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(ROOT, "include", "srt.h")]
    if not force and not _stale(LIB, srcs):
        return LIB
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, "srt_b200.cu"), "-o", LIB]
    subprocess.check_call(cmd)
    return LIB


HOST_LIB = os.path.join(HERE, "libsrt_host.so")


def build_host(force=False):
    """libsrt_host.so: the host-side helpers of include/srt_host.h (BVH builder, camera constructor); plain g++."""
    hdir = os.path.join(HERE, "host")
    srcs = [os.path.join(hdir, f) for f in ("srt_host.cpp", "srt_png.cpp", "srt_world.cpp", "srt_scenes.cpp")]
    deps = srcs + [os.path.join(hdir, "srt_json.hpp")] + [os.path.join(ROOT, "include", f) for f in ("srt_host.h", "srt.h", "srt_world.hpp")]
    if not force and not _stale(HOST_LIB, deps):
        return HOST_LIB
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-Wall",
                           "-I" + os.path.join(ROOT, "include")] + srcs + ["-o", HOST_LIB, "-lpthread", "-lz", "-ldl"])
    return HOST_LIB


EXAMPLE = os.path.join(ROOT, "examples", "sexy-raytracer")


def build_example(force=False):
    """examples/main.cpp: the reference's main.cpp over the host classes + the device (links libsrt_host.so only;
    libsrt_b200.so is dlopen'ed at run time)."""
    src = os.path.join(ROOT, "examples", "main.cpp")
    if not force and not _stale(EXAMPLE, [src, HOST_LIB, os.path.join(ROOT, "include", "srt_world.hpp")]):
        return EXAMPLE
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), src, "-L" + HERE, "-lsrt_host",
                           "-Wl,-rpath," + HERE, "-o", EXAMPLE])
    return EXAMPLE


def build_oracles():
    """The checkers: plain-C oracle always; the compiled reference only where /root/reference exists."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "all"])
