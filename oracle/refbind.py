"""ctypes bindings of the two CHECKERS (test infrastructure — never imported by the product):

  RefOracle  : oracle/_ref/libref_oracle.so, the reference's own CPU code (built from /root/reference)
  PortOracle : oracle/libsrt_oracle.so, the plain-C restatement (srt_oracle.c)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from sexy_raytracer_b200 import abi  # noqa: E402
from sexy_raytracer_b200.scene import Scene  # noqa: E402

REF_SO = os.path.join(_HERE, "_ref", "libref_oracle.so")
PORT_SO = os.path.join(_HERE, "libsrt_oracle.so")
REF_DATA = "/root/reference/data"

_vp = C.c_void_p


def _ptr(a):
    return a.ctypes.data_as(_vp)


def as_rays(o, d, time=None):
    n = len(o)
    r = np.zeros(n, abi.RAY_DTYPE)
    r["o"], r["dir"] = o, d
    if time is not None:
        r["time"] = time
    return r


class RefOracle:
    """The compiled reference.  Scene handles are opaque pointers owned by the library."""

    def __init__(self, path=REF_SO):
        self.lib = L = C.CDLL(path)
        L.ref_scene_native.restype = _vp
        L.ref_scene_native.argtypes = [C.c_int, C.c_char_p]
        L.ref_scene_from_desc.restype = _vp
        L.ref_scene_from_desc.argtypes = [C.POINTER(abi.SceneDesc)]
        L.ref_scene_free.argtypes = [_vp]
        L.ref_scene_sizes.argtypes = [_vp, _vp]
        L.ref_scene_export.argtypes = [_vp] * 17
        L.ref_trace_rays.argtypes = [_vp, _vp, C.c_int64, C.c_float, C.c_float, _vp]
        L.ref_count_tests.argtypes = [_vp, _vp, C.c_int64, C.c_float, C.c_float, _vp, _vp]
        L.ref_camera.argtypes = [_vp, C.POINTER(abi.Camera)]
        L.ref_render.restype = C.c_uint64
        L.ref_render.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_int, C.c_int, _vp, _vp]
        L.ref_resolve.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp]
        L.ref_getray_kat.argtypes = [_vp, _vp, C.c_int64, _vp]
        L.ref_shade_kat.argtypes = [_vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp]
        L.ref_texture_kat.argtypes = [_vp, C.c_int, _vp, C.c_int64, _vp]
        L.ref_aabb_kat.argtypes = [_vp, _vp, _vp, C.c_int64, _vp]
        L.ref_rng_seed.argtypes = [C.c_uint32]
        L.ref_rng_burn.argtypes = [C.c_uint64]
        L.ref_random_float.restype = C.c_float

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    # ---- scenes
    def native(self, which, data_dir=REF_DATA):
        h = self.lib.ref_scene_native(which, data_dir.encode())
        if not h:
            raise RuntimeError("ref_scene_native failed")
        return h

    def from_scene(self, scene):
        d = scene.desc()
        return self.lib.ref_scene_from_desc(C.byref(d))

    def free(self, h):
        self.lib.ref_scene_free(h)

    def export(self, h):
        """Flatten a reference scene into a Scene (texture bytes are copied)."""
        sz = np.zeros(5, np.int64)
        self.lib.ref_scene_sizes(h, _ptr(sz))
        nt, ns, nn, nm, nx = (int(x) for x in sz)
        s = Scene()
        f, i = np.float32, np.int32
        s.tri_v0, s.tri_v1, s.tri_v2 = (np.zeros((nt, 3), f) for _ in range(3))
        s.tri_uv0, s.tri_uv1, s.tri_uv2 = (np.zeros((nt, 2), f) for _ in range(3))
        s.tri_mat = np.zeros(nt, i)
        s.sph_c0, s.sph_c1 = np.zeros((ns, 3), f), np.zeros((ns, 3), f)
        s.sph_t0, s.sph_t1, s.sph_r = np.zeros(ns, f), np.zeros(ns, f), np.zeros(ns, f)
        s.sph_mat = np.zeros(ns, i)
        s.nodes = np.zeros(nn, abi.NODE_DTYPE)
        mats = (abi.Material * max(nm, 1))()
        texs = (abi.Texture * max(nx, 1))()
        self.lib.ref_scene_export(h, _ptr(s.tri_v0), _ptr(s.tri_v1), _ptr(s.tri_v2), _ptr(s.tri_uv0), _ptr(s.tri_uv1),
                                  _ptr(s.tri_uv2), _ptr(s.tri_mat), _ptr(s.sph_c0), _ptr(s.sph_c1), _ptr(s.sph_t0),
                                  _ptr(s.sph_t1), _ptr(s.sph_r), _ptr(s.sph_mat), _ptr(s.nodes),
                                  C.cast(mats, _vp), C.cast(texs, _vp))
        s.mat_kind = np.array([mats[k].kind for k in range(nm)], i)
        s.mat_tex = np.array([[mats[k].albedo_tex, mats[k].normal_tex, mats[k].metallic_tex, mats[k].roughness_tex]
                              for k in range(nm)], i).reshape(nm, 4)
        s.mat_albedo = np.array([list(mats[k].albedo) for k in range(nm)], f).reshape(nm, 4)
        s.mat_params = np.array([[mats[k].metalness, mats[k].roughness, mats[k].fuzz, mats[k].ir]
                                 for k in range(nm)], f).reshape(nm, 4)
        s.tex_kind = np.array([texs[k].kind for k in range(nx)], i)
        s.tex_color = np.array([list(texs[k].color) for k in range(nx)], f).reshape(nx, 3)
        s.tex_evenodd = np.array([[texs[k].even, texs[k].odd] for k in range(nx)], i).reshape(nx, 2)
        s.tex_whb = np.array([[texs[k].width, texs[k].height, texs[k].bpp] for k in range(nx)], i).reshape(nx, 3)
        s.tex_missing = np.zeros(nx, i)
        for k in range(nx):
            if texs[k].kind == abi.SRT_TEX_IMAGE:
                if texs[k].data:
                    n = texs[k].width * texs[k].height * texs[k].bpp
                    s.tex_data[k] = np.ctypeslib.as_array(texs[k].data, shape=(n,)).copy()
                else:
                    s.tex_missing[k] = 1
        return s

    # ---- the path
    def camera(self, cam_args):
        a = np.ascontiguousarray(cam_args, np.float32)
        cam = abi.Camera()
        self.lib.ref_camera(_ptr(a), C.byref(cam))
        return np.frombuffer(bytes(cam), np.float32).copy()

    def trace(self, h, rays, t_min=0.001, t_max=float("inf")):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.zeros(len(rays), abi.HIT_DTYPE)
        self.lib.ref_trace_rays(h, _ptr(rays), len(rays), t_min, t_max, _ptr(hits))
        return hits

    def count_tests(self, h, rays, t_min=0.001, t_max=float("inf")):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        b, l = C.c_uint64(), C.c_uint64()
        self.lib.ref_count_tests(h, _ptr(rays), len(rays), t_min, t_max, C.byref(b), C.byref(l))
        return b.value, l.value

    def render(self, h, cam_args, W, H, spp, depth, background, y0=0, y1=None, seed=None, want_rgba=True):
        """-> (sum [H,W,3] f32, rgba [H,W,4] u8 or None, segments).  seed=None keeps the engine running."""
        if seed is not None:
            self.lib.ref_rng_seed(seed)
        y1 = H if y1 is None else y1
        a = np.ascontiguousarray(cam_args, np.float32)
        bg = np.ascontiguousarray(background, np.float32)
        acc = np.zeros((H, W, 3), np.float32)
        rgba = np.zeros((H, W, 4), np.uint8) if want_rgba else None
        seg = self.lib.ref_render(h, _ptr(a), W, H, spp, depth, _ptr(bg), y0, y1, _ptr(acc),
                                  _ptr(rgba) if want_rgba else None)
        return acc, rgba, int(seg)

    def resolve(self, acc, spp):
        H, W, _ = acc.shape
        acc = np.ascontiguousarray(acc, np.float32)
        out = np.zeros((H, W, 4), np.uint8)
        self.lib.ref_resolve(_ptr(acc), W, H, spp, _ptr(out))
        return out

    # ---- known-answer generators
    def getray_kat(self, cam_args, inp):
        a = np.ascontiguousarray(cam_args, np.float32)
        inp = np.ascontiguousarray(inp, np.float32).reshape(-1, 5)
        out = np.zeros(len(inp), abi.RAY_DTYPE)
        self.lib.ref_getray_kat(_ptr(a), _ptr(inp), len(inp), _ptr(out))
        return out

    def shade_kat(self, h, rays, hits, xi):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.ascontiguousarray(hits, abi.HIT_DTYPE)
        xi = np.ascontiguousarray(xi, np.float32).reshape(-1, 4)
        n = len(rays)
        att, emit = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        out, sc = np.zeros(n, abi.RAY_DTYPE), np.zeros(n, np.int32)
        self.lib.ref_shade_kat(h, _ptr(rays), _ptr(hits), _ptr(xi), n, _ptr(att), _ptr(emit), _ptr(out), _ptr(sc))
        return att, emit, out, sc

    def texture_kat(self, h, tex, uvp):
        uvp = np.ascontiguousarray(uvp, np.float32).reshape(-1, 5)
        out = np.zeros((len(uvp), 3), np.float32)
        self.lib.ref_texture_kat(h, tex, _ptr(uvp), len(uvp), _ptr(out))
        return out

    def rotated_grid(self, s):
        out = np.zeros(2, np.float32)
        self.lib.ref_rotated_grid(C.c_int(int(s)), _ptr(out))
        return out

    def aabb_kat(self, rays, boxes, tminmax):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        boxes = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
        tmm = np.ascontiguousarray(tminmax, np.float32).reshape(-1, 2)
        out = np.zeros(len(rays), np.int32)
        self.lib.ref_aabb_kat(_ptr(rays), _ptr(boxes), _ptr(tmm), len(rays), _ptr(out))
        return out


class PortOracle:
    """The plain-C restatement (oracle/srt_oracle.c)."""

    def __init__(self, path=PORT_SO):
        self.lib = L = C.CDLL(path)
        L.so_scene_create.restype = _vp
        L.so_scene_create.argtypes = [C.POINTER(abi.SceneDesc)]
        L.so_scene_destroy.argtypes = [_vp]
        L.so_trace_rays.argtypes = [_vp, _vp, C.c_int64, C.c_float, C.c_float, _vp]
        L.so_count_tests.argtypes = [_vp, _vp, C.c_int64, C.c_float, C.c_float, _vp, _vp]
        L.so_aabb_kat.argtypes = [_vp, _vp, _vp, C.c_int64, _vp]
        L.so_texture_kat.argtypes = [_vp, C.c_int, _vp, C.c_int64, _vp]
        L.so_shade_kat.argtypes = [_vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp]
        L.so_getray_kat.argtypes = [C.POINTER(abi.Camera), _vp, C.c_int64, _vp]
        L.so_philox.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _vp]
        L.so_render.argtypes = [_vp, C.POINTER(abi.Camera), C.POINTER(abi.RenderParams), C.c_int, C.c_int, C.c_int,
                                _vp, C.POINTER(abi.Stats)]
        L.so_resolve.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp]

    @staticmethod
    def available():
        return os.path.exists(PORT_SO)

    def from_scene(self, scene):
        d = scene.desc()
        return self.lib.so_scene_create(C.byref(d))

    def free(self, h):
        self.lib.so_scene_destroy(h)

    def trace(self, h, rays, t_min=0.001, t_max=float("inf")):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.zeros(len(rays), abi.HIT_DTYPE)
        self.lib.so_trace_rays(h, _ptr(rays), len(rays), t_min, t_max, _ptr(hits))
        return hits

    def count_tests(self, h, rays, t_min=0.001, t_max=float("inf")):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        b, l = C.c_uint64(), C.c_uint64()
        self.lib.so_count_tests(h, _ptr(rays), len(rays), t_min, t_max, C.byref(b), C.byref(l))
        return b.value, l.value

    def aabb_kat(self, rays, boxes, tminmax):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        boxes = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
        tmm = np.ascontiguousarray(tminmax, np.float32).reshape(-1, 2)
        out = np.zeros(len(rays), np.int32)
        self.lib.so_aabb_kat(_ptr(rays), _ptr(boxes), _ptr(tmm), len(rays), _ptr(out))
        return out

    def texture_kat(self, h, tex, uvp):
        uvp = np.ascontiguousarray(uvp, np.float32).reshape(-1, 5)
        out = np.zeros((len(uvp), 3), np.float32)
        self.lib.so_texture_kat(h, tex, _ptr(uvp), len(uvp), _ptr(out))
        return out

    def shade_kat(self, h, rays, hits, rvec_xi):
        rays = np.ascontiguousarray(rays, abi.RAY_DTYPE)
        hits = np.ascontiguousarray(hits, abi.HIT_DTYPE)
        xi = np.ascontiguousarray(rvec_xi, np.float32).reshape(-1, 4)
        n = len(rays)
        att, emit = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        out, sc = np.zeros(n, abi.RAY_DTYPE), np.zeros(n, np.int32)
        self.lib.so_shade_kat(h, _ptr(rays), _ptr(hits), _ptr(xi), n, _ptr(att), _ptr(emit), _ptr(out), _ptr(sc))
        return att, emit, out, sc

    def getray_kat(self, camera24, inp):
        cam = abi.Camera()
        C.memmove(C.byref(cam), np.ascontiguousarray(camera24, np.float32).ctypes.data, C.sizeof(cam))
        inp = np.ascontiguousarray(inp, np.float32).reshape(-1, 5)
        out = np.zeros(len(inp), abi.RAY_DTYPE)
        self.lib.so_getray_kat(C.byref(cam), _ptr(inp), len(inp), _ptr(out))
        return out

    def rotated_grid(self, s):
        x, y = C.c_float(), C.c_float()
        self.lib.so_rotated_grid(C.c_int(int(s)), C.byref(x), C.byref(y))
        return np.array([x.value, y.value], np.float32)

    def philox(self, seed, pixel, sample, block):
        out = np.zeros(4, np.uint32)
        self.lib.so_philox(seed, pixel, sample, block, _ptr(out))
        return out

    def render(self, h, scene, params=None, y0=0, y1=None, nthreads=1):
        """-> (sum [H,W,3] f32, Stats)"""
        P = params if params is not None else scene.render_params()
        cam = scene.camera_struct()
        y1 = P.height if y1 is None else y1
        acc = np.zeros((P.height, P.width, 3), np.float32)
        st = abi.Stats()
        self.lib.so_render(h, C.byref(cam), C.byref(P), y0, y1, nthreads, _ptr(acc), C.byref(st))
        return acc, st

    def resolve(self, acc, spp):
        H, W, _ = acc.shape
        acc = np.ascontiguousarray(acc, np.float32)
        out = np.zeros((H, W, 4), np.uint8)
        self.lib.so_resolve(_ptr(acc), W, H, spp, _ptr(out))
        return out
