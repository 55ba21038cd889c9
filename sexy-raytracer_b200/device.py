"""Thin ctypes binding of libsrt_b200.so (include/srt.h).  No arithmetic happens in Python and there
is no fallback: if the CUDA library is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SRT_B200_LIBRARY") or os.path.join(_HERE, "libsrt_b200.so")   # the override is for A/B builds (tools/)
_lib = None


class SrtError(RuntimeError):
    pass


def lib():
    """The loaded C-ABI library (loaded once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SrtError("libsrt_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
        _lib = abi.bind(C.CDLL(LIB_PATH))
    return _lib


def _check(rc):
    if rc != 0:
        L = lib()
        raise SrtError("%s: %s" % (L.srt_status_string(rc).decode(), L.srt_last_error().decode()))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Device:
    """srt_device: one GPU.  Replaces glDevice::init/terminate (gl.h:28,31)."""

    def __init__(self, gpu_id=0):
        self.h = C.c_void_p()
        _check(lib().srt_device_create(gpu_id, C.byref(self.h)))
        self.gpu_id = gpu_id

    def close(self):
        if self.h:
            lib().srt_device_destroy(self.h)
            self.h = C.c_void_p()

    def set_counters(self, on):
        _check(lib().srt_set_counters(self.h, 1 if on else 0))

    def last_counters(self):
        out = np.zeros(5, np.uint64)
        _check(lib().srt_last_counters(self.h, _ptr(out)))
        return {"node_visits": int(out[0]), "leaf_tests": int(out[1]), "exact_box_tests": int(out[2]),
                "sphere_leaf_tests": int(out[3]), "triangle_hits": int(out[4])}

    @property
    def stream(self):
        return lib().srt_device_stream(self.h)

    def set_stream(self, cuda_stream_ptr):
        """Launch on the caller's CUDA stream (an int / void*; None restores the device's own stream)."""
        _check(lib().srt_device_set_stream(self.h, cuda_stream_ptr))

    def upload(self, scene):
        return DeviceScene(self, scene)

    def resolve_dev(self, sum_dev_ptr, W, H, spp, rgba_dev_ptr):
        _check(lib().srt_resolve(self.h, sum_dev_ptr, W, H, spp, rgba_dev_ptr))

    def aabb_kat(self, rays, boxes, tminmax, predicate=0):
        """-> (decision [n] i32, path [n] i32); see srt_aabb_kat in include/srt.h"""
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        boxes = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
        tmm = np.ascontiguousarray(tminmax, np.float32).reshape(-1, 2)
        dec, path = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.int32)
        _check(lib().srt_aabb_kat(self.h, _ptr(rays), _ptr(boxes), _ptr(tmm), len(rays), predicate, _ptr(dec), _ptr(path)))
        return dec, path

    def getray_kat(self, camera, inp):
        inp = np.ascontiguousarray(inp, np.float32).reshape(-1, 5)
        out = np.zeros(len(inp), abi.RAY_DTYPE)
        _check(lib().srt_getray_kat(self.h, C.byref(camera), _ptr(inp), len(inp), _ptr(out)))
        return out

    def primary_rays(self, camera, params, pixel_xy, sample):
        pixel_xy = np.ascontiguousarray(pixel_xy, np.int32).reshape(-1, 2)
        sample = np.ascontiguousarray(sample, np.int32)
        out = np.zeros(len(sample), abi.RAY_DTYPE)
        _check(lib().srt_primary_rays(self.h, C.byref(camera), C.byref(params), _ptr(pixel_xy), _ptr(sample), len(sample),
                                      _ptr(out)))
        return out


class DeviceScene:
    """srt_scene: a flattened scene resident on the GPU."""

    def __init__(self, device, scene):
        self.device, self.scene = device, scene
        self.h = C.c_void_p()
        d = scene.desc()
        _check(lib().srt_scene_create(device.h, C.byref(d), C.byref(self.h)))

    def close(self):
        if self.h:
            lib().srt_scene_destroy(self.h)
            self.h = C.c_void_p()

    @property
    def device_bytes(self):
        return int(lib().srt_scene_device_bytes(self.h))

    def trace(self, rays, t_min=0.001, t_max=float("inf")):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.zeros(len(rays), abi.HIT_DTYPE)
        _check(lib().srt_trace_rays(self.h, _ptr(rays), len(rays), t_min, t_max, _ptr(hits)))
        return hits

    def trace_wavefront(self, rays):
        """srt_trace_rays_wavefront: one production EXTEND wave over explicit rays -> (prim [n] i32, t [n] f32)"""
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        prim, t = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.float32)
        _check(lib().srt_trace_rays_wavefront(self.h, _ptr(rays), len(rays), _ptr(prim), _ptr(t)))
        return prim, t

    def trace_wavefront_static(self, rays_dev_ptr, n, out_prim_dev_ptr, out_t_dev_ptr):
        """srt_trace_rays_wavefront_static: the same wave on DEVICE pointers (ints / void*), allocating nothing; the caller owns
        the three device buffers.  At most srt_static_max_rays() rays per call; one call at a time per GPU."""
        _check(lib().srt_trace_rays_wavefront_static(self.h, rays_dev_ptr, n, out_prim_dev_ptr, out_t_dev_ptr))

    def render(self, params=None, want_sum=True, want_rgba=True, camera=None):
        """srt_render with host buffers -> (sum [H,W,3] f32 | None, rgba [H,W,4] u8 | None, Stats)"""
        P = params if params is not None else self.scene.render_params()
        cam = camera if camera is not None else self.scene.camera_struct()
        acc = np.zeros((P.height, P.width, 3), np.float32) if want_sum else None
        rgba = np.zeros((P.height, P.width, 4), np.uint8) if want_rgba else None
        st = abi.Stats()
        _check(lib().srt_render(self.h, C.byref(cam), C.byref(P), _ptr(acc) if want_sum else None,
                                _ptr(rgba) if want_rgba else None, C.byref(st)))
        return acc, rgba, st

    def render_accumulate(self, accum_dev_ptr, params=None, camera=None):
        """Adds this call's samples into a device buffer [H*W*3] f32 (e.g. a torch tensor's data_ptr())."""
        P = params if params is not None else self.scene.render_params()
        cam = camera if camera is not None else self.scene.camera_struct()
        st = abi.Stats()
        _check(lib().srt_render_accumulate(self.h, C.byref(cam), C.byref(P), accum_dev_ptr, C.byref(st)))
        return st

    def shade_kat(self, rays, hits, rvec_xi):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.ascontiguousarray(hits, abi.HIT_DTYPE)
        xi = np.ascontiguousarray(rvec_xi, np.float32).reshape(-1, 4)
        n = len(rays)
        att, emit = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        out, sc = np.zeros(n, abi.RAY_DTYPE), np.zeros(n, np.int32)
        _check(lib().srt_shade_kat(self.h, _ptr(rays), _ptr(hits), _ptr(xi), n, _ptr(att), _ptr(emit), _ptr(out), _ptr(sc)))
        return att, emit, out, sc

    def texture_kat(self, tex, uvp):
        uvp = np.ascontiguousarray(uvp, np.float32).reshape(-1, 5)
        out = np.zeros((len(uvp), 3), np.float32)
        _check(lib().srt_texture_kat(self.h, tex, _ptr(uvp), len(uvp), _ptr(out)))
        return out


class DeviceGroup:
    """srt_device_group: several GPUs behind one handle (gpu_ids=None: every visible GPU)."""

    def __init__(self, gpu_ids=None):
        self.h = C.c_void_p()
        if gpu_ids is None:
            _check(lib().srt_device_group_create(None, 0, C.byref(self.h)))
        else:
            ids = (C.c_int * len(gpu_ids))(*gpu_ids)
            _check(lib().srt_device_group_create(ids, len(gpu_ids), C.byref(self.h)))

    def __len__(self):
        return int(lib().srt_device_group_size(self.h))

    def close(self):
        if self.h:
            lib().srt_device_group_destroy(self.h)
            self.h = C.c_void_p()

    def upload(self, scene):
        return SceneGroup(self, scene)


class SceneGroup:
    """srt_scene_group: one replica of a flattened scene per GPU of the group."""

    def __init__(self, group, scene):
        self.group, self.scene = group, scene
        self.h = C.c_void_p()
        d = scene.desc()
        _check(lib().srt_scene_group_create(group.h, C.byref(d), C.byref(self.h)))

    def close(self):
        if self.h:
            lib().srt_scene_group_destroy(self.h)
            self.h = C.c_void_p()

    def render(self, params=None, want_sum=True, want_rgba=True, camera=None):
        """srt_render_group -> (sum | None, rgba | None, Stats)"""
        P = params if params is not None else self.scene.render_params()
        cam = camera if camera is not None else self.scene.camera_struct()
        acc = np.zeros((P.height, P.width, 3), np.float32) if want_sum else None
        rgba = np.zeros((P.height, P.width, 4), np.uint8) if want_rgba else None
        st = abi.Stats()
        _check(lib().srt_render_group(self.h, C.byref(cam), C.byref(P), _ptr(acc) if want_sum else None,
                                      _ptr(rgba) if want_rgba else None, C.byref(st)))
        return acc, rgba, st


class Film:
    """srt_film: progressive accumulation (the per-frame loop of glDevice::rtFrame, main.cpp:195-199, headless)."""

    def __init__(self, device, width, height):
        self.device, self.width, self.height = device, width, height
        self.h = C.c_void_p()
        _check(lib().srt_film_create(device.h, width, height, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().srt_film_destroy(self.h)
            self.h = C.c_void_p()

    @property
    def samples(self):
        return int(lib().srt_film_samples(self.h))

    def add_samples(self, dscene, count, params=None, camera=None):
        P = params if params is not None else dscene.scene.render_params(width=self.width, height=self.height)
        cam = camera if camera is not None else dscene.scene.camera_struct()
        st = abi.Stats()
        _check(lib().srt_film_add_samples(self.h, dscene.h, C.byref(cam), C.byref(P), count, C.byref(st)))
        return st

    def resolve(self, want_sum=True, want_rgba=True):
        acc = np.zeros((self.height, self.width, 3), np.float32) if want_sum else None
        rgba = np.zeros((self.height, self.width, 4), np.uint8) if want_rgba else None
        _check(lib().srt_film_resolve(self.h, _ptr(rgba) if want_rgba else None, _ptr(acc) if want_sum else None))
        return acc, rgba

    def load(self, sum_rgb, samples_done):
        a = np.ascontiguousarray(sum_rgb, np.float32)
        assert a.size == self.width * self.height * 3
        _check(lib().srt_film_load(self.h, _ptr(a), samples_done))
