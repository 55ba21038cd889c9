"""ctypes mirror of include/srt.h (the C ABI of libsrt_b200.so).

Only struct layouts and prototypes live here; no arithmetic.  Every struct below must stay
byte-identical to its C twin (tests/test_abi.py checks the sizes the library reports).
"""
import ctypes as C

SRT_BVH_NONE = -(2 ** 31)

SRT_TEX_SOLID, SRT_TEX_CHECKER, SRT_TEX_IMAGE = 0, 1, 2
SRT_MAT_PBR, SRT_MAT_METAL, SRT_MAT_DIELECTRIC, SRT_MAT_LIGHT = 0, 1, 2, 3

c_float_p = C.POINTER(C.c_float)
c_int32_p = C.POINTER(C.c_int32)
c_uint8_p = C.POINTER(C.c_uint8)


class BvhNode(C.Structure):
    _fields_ = [("bmin", C.c_float * 3), ("bmax", C.c_float * 3), ("left", C.c_int32), ("right", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", C.c_float * 3), ("even", C.c_int32), ("odd", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("bpp", C.c_int32), ("data", c_uint8_p)]


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("albedo_tex", C.c_int32), ("normal_tex", C.c_int32),
                ("metallic_tex", C.c_int32), ("roughness_tex", C.c_int32), ("albedo", C.c_float * 4),
                ("metalness", C.c_float), ("roughness", C.c_float), ("fuzz", C.c_float), ("ir", C.c_float),
                ("flags", C.c_int32)]


MATF_PACKED_MR = 1


class SceneDesc(C.Structure):
    _fields_ = [("n_tri", C.c_int64),
                ("tri_v0", c_float_p), ("tri_v1", c_float_p), ("tri_v2", c_float_p),
                ("tri_uv0", c_float_p), ("tri_uv1", c_float_p), ("tri_uv2", c_float_p),
                ("tri_mat", c_int32_p),
                ("n_sph", C.c_int64),
                ("sph_c0", c_float_p), ("sph_c1", c_float_p), ("sph_t0", c_float_p), ("sph_t1", c_float_p),
                ("sph_r", c_float_p), ("sph_mat", c_int32_p),
                ("n_nodes", C.c_int64), ("nodes", C.POINTER(BvhNode)),
                ("n_mat", C.c_int32), ("materials", C.POINTER(Material)),
                ("n_tex", C.c_int32), ("textures", C.POINTER(Texture)),
                ("tri_n0", c_float_p), ("tri_n1", c_float_p), ("tri_n2", c_float_p),
                ("tri_tan0", c_float_p), ("tri_tan1", c_float_p), ("tri_tan2", c_float_p)]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("lleft", C.c_float * 3), ("horizontal", C.c_float * 3),
                ("vertical", C.c_float * 3), ("w", C.c_float * 3), ("hor", C.c_float * 3), ("vert", C.c_float * 3),
                ("lens_radius", C.c_float), ("t0", C.c_float), ("t1", C.c_float)]


class Ray(C.Structure):
    _fields_ = [("o", C.c_float * 3), ("dir", C.c_float * 3), ("time", C.c_float)]


class Hit(C.Structure):
    _fields_ = [("prim", C.c_int32), ("mat", C.c_int32), ("front_face", C.c_int32), ("t", C.c_float),
                ("p", C.c_float * 3), ("normal", C.c_float * 3), ("tangent", C.c_float * 3),
                ("bitangent", C.c_float * 3), ("uv", C.c_float * 2)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("sample_begin", C.c_int32),
                ("sample_count", C.c_int32), ("max_bounce", C.c_int32), ("background", C.c_float * 3),
                ("seed", C.c_uint64), ("rr_begin", C.c_int32), ("variant", C.c_int32), ("sampler", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("segments", C.c_uint64), ("node_visits", C.c_uint64),
                ("leaf_tests", C.c_uint64), ("kernel_ms", C.c_float), ("launches", C.c_int32),
                ("extend_ms", C.c_float), ("shade_ms", C.c_float), ("extend_launches", C.c_int32), ("variant", C.c_int32),
                ("extend_kernel", C.c_int32), ("fused", C.c_int32)]


# numpy dtypes with the same layout (for bulk arrays of rays / hits / nodes)
import numpy as np  # noqa: E402

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("dir", "<f4", 3), ("time", "<f4")])
HIT_DTYPE = np.dtype([("prim", "<i4"), ("mat", "<i4"), ("front_face", "<i4"), ("t", "<f4"), ("p", "<f4", 3),
                      ("normal", "<f4", 3), ("tangent", "<f4", 3), ("bitangent", "<f4", 3), ("uv", "<f4", 2)])
NODE_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("left", "<i4"), ("right", "<i4")])
assert RAY_DTYPE.itemsize == C.sizeof(Ray) == 28
assert HIT_DTYPE.itemsize == C.sizeof(Hit) == 72
assert NODE_DTYPE.itemsize == C.sizeof(BvhNode) == 32

# every symbol include/srt.h declares: name -> (restype, argtypes)
_vp = C.c_void_p
PROTOTYPES = {
    "srt_device_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "srt_device_destroy": (None, [_vp]),
    "srt_scene_create": (C.c_int, [_vp, C.POINTER(SceneDesc), C.POINTER(_vp)]),
    "srt_scene_destroy": (None, [_vp]),
    "srt_scene_device_bytes": (C.c_int64, [_vp]),
    "srt_trace_rays": (C.c_int, [_vp, _vp, C.c_int64, C.c_float, C.c_float, _vp]),
    "srt_trace_rays_wavefront": (C.c_int, [_vp, _vp, C.c_int64, _vp, _vp]),
    "srt_trace_rays_wavefront_static": (C.c_int, [_vp, _vp, C.c_int64, _vp, _vp]),
    "srt_static_max_rays": (C.c_int64, []),
    "srt_aabb_kat": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int64, C.c_int32, _vp, _vp]),
    "srt_render": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp, C.POINTER(Stats)]),
    "srt_render_accumulate": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, C.POINTER(Stats)]),
    "srt_resolve": (C.c_int, [_vp, _vp, C.c_int32, C.c_int32, C.c_int32, _vp]),
    "srt_device_group_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]),
    "srt_device_group_destroy": (None, [_vp]),
    "srt_device_group_size": (C.c_int32, [_vp]),
    "srt_device_group_device": (_vp, [_vp, C.c_int32]),
    "srt_scene_group_create": (C.c_int, [_vp, C.POINTER(SceneDesc), C.POINTER(_vp)]),
    "srt_scene_group_destroy": (None, [_vp]),
    "srt_scene_group_scene": (_vp, [_vp, C.c_int32]),
    "srt_sample_split": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "srt_render_group": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp, C.POINTER(Stats)]),
    "srt_film_create": (C.c_int, [_vp, C.c_int32, C.c_int32, C.POINTER(_vp)]),
    "srt_film_destroy": (None, [_vp]),
    "srt_film_add_samples": (C.c_int, [_vp, _vp, C.POINTER(Camera), C.POINTER(RenderParams), C.c_int32, C.POINTER(Stats)]),
    "srt_film_samples": (C.c_int32, [_vp]),
    "srt_film_resolve": (C.c_int, [_vp, _vp, _vp]),
    "srt_film_load": (C.c_int, [_vp, _vp, C.c_int32]),
    "srt_shade_kat": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp]),
    "srt_texture_kat": (C.c_int, [_vp, C.c_int32, _vp, C.c_int64, _vp]),
    "srt_getray_kat": (C.c_int, [_vp, C.POINTER(Camera), _vp, C.c_int64, _vp]),
    "srt_primary_rays": (C.c_int, [_vp, C.POINTER(Camera), C.POINTER(RenderParams), _vp, _vp, C.c_int64, _vp]),
    "srt_last_counters": (C.c_int, [_vp, _vp]),
    "srt_abi_version": (C.c_int, []),
    "srt_status_string": (C.c_char_p, [C.c_int]),
    "srt_last_error": (C.c_char_p, []),
    "srt_set_counters": (C.c_int, [_vp, C.c_int]),
    "srt_device_stream": (_vp, [_vp]),
    "srt_device_set_stream": (C.c_int, [_vp, _vp]),
}


def bind(lib):
    """Attach restype/argtypes of every declared entry point; raises AttributeError if one is missing."""
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
