"""ctypes binding of libsrt_host.so (include/srt_host.h): the host-side steps immediately before the path whose results
must equal the reference's - BVH construction (bvh.h:55-95), primitive boxes, the camera constructor (camera.h:10-38)."""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsrt_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from . import build
            build.build_host()
        L = C.CDLL(LIB_PATH)
        L.srt_host_last_error.restype = C.c_char_p
        L.srt_bvh_build.argtypes = [C.POINTER(abi.SceneDesc), C.c_float, C.c_float, C.c_uint64, C.c_int,
                                    C.POINTER(C.POINTER(abi.BvhNode)), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        L.srt_bvh_free.argtypes = [C.POINTER(abi.BvhNode)]
        L.srt_prim_boxes.argtypes = [C.POINTER(abi.SceneDesc), C.c_float, C.c_float, C.c_void_p]
        L.srt_camera_make.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_float] * 6 + [C.POINTER(abi.Camera)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("srt_host: status %d: %s" % (rc, lib().srt_host_last_error().decode()))


def build_bvh(scene, time0=0.0, time1=1.0, rng_skip=0, threads=0):
    """bvhNode(objects, 0, n, time0, time1) over all primitives of `scene` -> (nodes [NODE_DTYPE], draws consumed)."""
    d = scene.desc()
    out = C.POINTER(abi.BvhNode)()
    n = C.c_int64()
    draws = C.c_uint64()
    _check(lib().srt_bvh_build(C.byref(d), time0, time1, rng_skip, threads, C.byref(out), C.byref(n), C.byref(draws)))
    if n.value == 0:
        return np.zeros(0, abi.NODE_DTYPE), 0
    buf = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_uint8)), shape=(n.value * 32,))
    nodes = buf.view(abi.NODE_DTYPE).copy()
    lib().srt_bvh_free(out)
    return nodes, int(draws.value)


def prim_boxes(scene, time0=0.0, time1=1.0):
    d = scene.desc()
    out = np.zeros((scene.n_prim, 6), np.float32)
    _check(lib().srt_prim_boxes(C.byref(d), time0, time1, out.ctypes.data_as(C.c_void_p)))
    return out


def make_camera(eye, look_at, up, vfov, aspect, aperture, focus_dist, t0=0.0, t1=1.0):
    """-> flat float32[24] in srt_camera layout (what Scene.camera holds)."""
    e, l, u = (np.ascontiguousarray(x, np.float32) for x in (eye, look_at, up))
    cam = abi.Camera()
    _check(lib().srt_camera_make(e.ctypes.data, l.ctypes.data, u.ctypes.data, vfov, aspect, aperture, focus_dist, t0, t1,
                                 C.byref(cam)))
    return np.frombuffer(bytes(cam), np.float32).copy()
