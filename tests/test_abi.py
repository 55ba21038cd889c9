"""CPU-side checks of the drop-in boundary: libsrt_b200.so loads without a GPU and exports every symbol include/srt.h
declares; struct layouts of the ctypes mirror match the header's; compute calls fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT
from sexy_raytracer_b200 import abi, build


@pytest.fixture(scope="module")
def lib():
    build.build_cuda()                      # nvcc cross-compiles sm_100a without a GPU
    return C.CDLL(build.LIB)


def declared_symbols(header="srt.h"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z_0-9]+)\s*\(", src)))


def test_abi_symbols(lib):
    names = declared_symbols()
    assert len(names) >= 20 and "srt_render" in names and "srt_trace_rays" in names
    for n in names:
        assert hasattr(lib, n), "libsrt_b200.so does not export %s" % n
    lib.srt_abi_version.restype = C.c_int
    assert lib.srt_abi_version() == 1


def test_host_abi_symbols():
    """libsrt_host.so (plain C++, no CUDA) exports every symbol include/srt_host.h declares, and the C++ classes of
    include/srt_world.hpp link against it (examples/main.cpp is the link test)."""
    build.build_host()
    L = C.CDLL(build.HOST_LIB)
    names = declared_symbols("srt_host.h")
    assert len(names) >= 15 and "srt_bvh_build" in names and "srt_png_write" in names and "srt_checkpoint_read" in names
    for n in names:
        assert hasattr(L, n), "libsrt_host.so does not export %s" % n
    assert os.access(build.build_example(), os.X_OK)


def test_struct_sizes_match_header():
    assert C.sizeof(abi.BvhNode) == 32 and abi.NODE_DTYPE.itemsize == 32
    assert C.sizeof(abi.Ray) == 28 and C.sizeof(abi.Camera) == 24 * 4
    assert C.sizeof(abi.Hit) == abi.HIT_DTYPE.itemsize == 4 * 4 + 12 * 4 + 2 * 4
    assert C.sizeof(abi.Stats) == 4 * 8 + 8 * 4


def test_no_cpu_fallback(lib):
    """Without a CUDA device every compute entry point must fail, not compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = abi.bind(lib)
    h = C.c_void_p()
    rc = L.srt_device_create(0, C.byref(h))
    assert rc != 0 and not h
    assert L.srt_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sexy-raytracer_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(d, f)).read()
                assert "oracle" not in txt.replace("# noqa", "") or f == "build.py", "%s mentions oracle/" % f
