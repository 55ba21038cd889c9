"""sexy-raytracer_b200 — B200-native path-tracing device behind the reference's scene/camera surface.

Holds only what the hot path needs: csrc/ (CUDA kernels + the C ABI of include/srt.h, built into
libsrt_b200.so), host/ (C++ host classes mirroring the reference's camera/hittable/material/texture
API and owning the flattening), and thin ctypes bindings for tests and bench.
"""
from . import abi  # noqa: F401
from .scene import Scene  # noqa: F401
