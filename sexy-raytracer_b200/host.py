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


# ---- whole scenes through the C++ host classes (include/srt_world.hpp) ----------------------------------------------
def _bind_world(L):
    if getattr(L, "_world_bound", False):
        return L
    L.srt_host_scene_reference.argtypes = [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]
    L.srt_host_scene_gltf.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
    L.srt_host_scene_desc.argtypes = [C.c_void_p]
    L.srt_host_scene_desc.restype = C.POINTER(abi.SceneDesc)
    L.srt_host_scene_draws.argtypes = [C.c_void_p]
    L.srt_host_scene_draws.restype = C.c_uint64
    L.srt_host_scene_free.argtypes = [C.c_void_p]
    L.srt_png_read.argtypes = [C.c_char_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.POINTER(C.c_uint8))]
    L.srt_png_read_memory.argtypes = [C.c_char_p, C.c_int64, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                      C.POINTER(C.POINTER(C.c_uint8))]
    L.srt_png_write.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32]
    L.srt_host_free.argtypes = [C.c_void_p]
    L.srt_checkpoint_write.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.c_void_p]
    L.srt_checkpoint_read.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_uint64),
                                      C.POINTER(C.POINTER(C.c_float))]
    L.srt_bvh_build_ordered.argtypes = [C.POINTER(abi.SceneDesc), C.c_void_p, C.c_float, C.c_float, C.c_uint64, C.c_int,
                                        C.POINTER(C.POINTER(abi.BvhNode)), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
    L._world_bound = True
    return L


def scene_from_desc(d):
    """Copies everything an srt_scene_desc points at into a Scene (numpy arrays)."""
    from .scene import Scene

    def arr(p, n, cols, dt):
        if n == 0:
            return np.zeros((0, cols) if cols > 1 else 0, dt)
        a = np.ctypeslib.as_array(p, shape=(n * cols,)).copy()
        return a.reshape(n, cols) if cols > 1 else a

    s = Scene()
    nt, ns = int(d.n_tri), int(d.n_sph)
    s.tri_v0, s.tri_v1, s.tri_v2 = (arr(p, nt, 3, np.float32) for p in (d.tri_v0, d.tri_v1, d.tri_v2))
    s.tri_uv0, s.tri_uv1, s.tri_uv2 = (arr(p, nt, 2, np.float32) for p in (d.tri_uv0, d.tri_uv1, d.tri_uv2))
    s.tri_mat = arr(d.tri_mat, nt, 1, np.int32)
    if nt and d.tri_n0:
        s.tri_n0, s.tri_n1, s.tri_n2 = (arr(p, nt, 3, np.float32) for p in (d.tri_n0, d.tri_n1, d.tri_n2))
    if nt and d.tri_tan0:
        s.tri_tan0, s.tri_tan1, s.tri_tan2 = (arr(p, nt, 4, np.float32) for p in (d.tri_tan0, d.tri_tan1, d.tri_tan2))
    s.sph_c0, s.sph_c1 = arr(d.sph_c0, ns, 3, np.float32), arr(d.sph_c1, ns, 3, np.float32)
    s.sph_t0, s.sph_t1, s.sph_r = (arr(p, ns, 1, np.float32) for p in (d.sph_t0, d.sph_t1, d.sph_r))
    s.sph_mat = arr(d.sph_mat, ns, 1, np.int32)
    nn = int(d.n_nodes)
    if nn:
        s.nodes = np.ctypeslib.as_array(C.cast(d.nodes, C.POINTER(C.c_uint8)), shape=(nn * 32,)).view(abi.NODE_DTYPE).copy()
    nm, nx = int(d.n_mat), int(d.n_tex)
    s.mat_kind = np.array([d.materials[i].kind for i in range(nm)], np.int32)
    s.mat_tex = np.array([[d.materials[i].albedo_tex, d.materials[i].normal_tex, d.materials[i].metallic_tex,
                           d.materials[i].roughness_tex] for i in range(nm)], np.int32).reshape(nm, 4)
    s.mat_albedo = np.array([list(d.materials[i].albedo) for i in range(nm)], np.float32).reshape(nm, 4)
    s.mat_params = np.array([[d.materials[i].metalness, d.materials[i].roughness, d.materials[i].fuzz, d.materials[i].ir]
                             for i in range(nm)], np.float32).reshape(nm, 4)
    flags = np.array([d.materials[i].flags for i in range(nm)], np.int32)
    if flags.any():
        s.mat_flags = flags
    s.tex_kind = np.array([d.textures[i].kind for i in range(nx)], np.int32)
    s.tex_color = np.array([list(d.textures[i].color) for i in range(nx)], np.float32).reshape(nx, 3)
    s.tex_evenodd = np.array([[d.textures[i].even, d.textures[i].odd] for i in range(nx)], np.int32).reshape(nx, 2)
    s.tex_whb = np.array([[d.textures[i].width, d.textures[i].height, d.textures[i].bpp] for i in range(nx)], np.int32).reshape(nx, 3)
    s.tex_missing = np.zeros(nx, np.int32)
    for i in range(nx):
        t = d.textures[i]
        if t.kind == abi.SRT_TEX_IMAGE:
            if t.data:
                s.tex_data[i] = np.ctypeslib.as_array(t.data, shape=(t.width * t.height * t.bpp,)).copy()
            else:
                s.tex_missing[i] = 1
    return s


def reference_scene(which, data_dir):
    """The reference's scene `which` (1..4, include/srt_host.h) built by the C++ host classes -> (Scene, random draws)."""
    L = _bind_world(lib())
    h = C.c_void_p()
    _check(L.srt_host_scene_reference(which, os.fsencode(data_dir), C.byref(h)))
    try:
        return scene_from_desc(L.srt_host_scene_desc(h).contents), int(L.srt_host_scene_draws(h))
    finally:
        L.srt_host_scene_free(h)


def gltf_scene(path, texture_dir=None, reference_quirks=True, smooth_normals=False, sample_metallic_roughness=False, vertex_tangents=False):
    """A glTF / GLB file through model::create(...)->init() + bvhNode + flatScene -> Scene (triangles only).  The last two
    options are not reference behaviour (include/srt_world.hpp: gltfOptions)."""
    L = _bind_world(lib())
    h = C.c_void_p()
    opts = (1 if reference_quirks else 0) | (2 if smooth_normals else 0) | (4 if sample_metallic_roughness else 0) | (8 if vertex_tangents else 0)
    _check(L.srt_host_scene_gltf(os.fsencode(path), os.fsencode(texture_dir) if texture_dir else None, opts, C.byref(h)))
    try:
        return scene_from_desc(L.srt_host_scene_desc(h).contents)
    finally:
        L.srt_host_scene_free(h)


def png_read(path_or_bytes, req_comp):
    """stbi_load(path, &w, &h, &n, req_comp) -> uint8 array [h, w, req_comp]."""
    L = _bind_world(lib())
    w, h, p = C.c_int32(), C.c_int32(), C.POINTER(C.c_uint8)()
    if isinstance(path_or_bytes, (bytes, bytearray)):
        _check(L.srt_png_read_memory(bytes(path_or_bytes), len(path_or_bytes), req_comp, C.byref(w), C.byref(h), C.byref(p)))
    else:
        _check(L.srt_png_read(os.fsencode(path_or_bytes), req_comp, C.byref(w), C.byref(h), C.byref(p)))
    a = np.ctypeslib.as_array(p, shape=(h.value * w.value * req_comp,)).copy().reshape(h.value, w.value, req_comp)
    L.srt_host_free(p)
    return a


def png_write(path, pixels):
    """stbi_write_png(path, w, h, comp, data, w*comp) for a uint8 array [h, w, comp]."""
    L = _bind_world(lib())
    a = np.ascontiguousarray(pixels, np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    _check(L.srt_png_write(os.fsencode(path), a.shape[1], a.shape[0], a.shape[2], a.ctypes.data_as(C.c_void_p), 0))


def pfm_write(path, rgb, scale=1.0):
    """srt_pfm_write: fp32 radiance [h, w, 3], row 0 = top, as a little-endian colour PFM (values times `scale`)."""
    L = _bind_world(lib())
    a = np.ascontiguousarray(rgb, np.float32)
    L.srt_pfm_write.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_void_p, C.c_float]
    _check(L.srt_pfm_write(os.fsencode(path), a.shape[1], a.shape[0], a.ctypes.data_as(C.c_void_p), scale))


def pfm_read(path):
    """srt_pfm_read -> float32 [h, w, 3], row 0 = top."""
    L = _bind_world(lib())
    w, h, p = C.c_int32(), C.c_int32(), C.c_void_p()
    L.srt_pfm_read.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_void_p)]
    _check(L.srt_pfm_read(os.fsencode(path), C.byref(w), C.byref(h), C.byref(p)))
    try:
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), (h.value, w.value, 3)).copy()
    finally:
        L.srt_host_free(p)


def checkpoint_write(path, sum_rgb, samples_done, seed=1):
    """srt_checkpoint_write: the film of a progressive render (fp32 sums [h, w, 3] + samples per pixel so far)."""
    a = np.ascontiguousarray(sum_rgb, np.float32)
    _check(_bind_world(lib()).srt_checkpoint_write(os.fsencode(path), a.shape[1], a.shape[0], samples_done, seed, a.ctypes.data_as(C.c_void_p)))


def checkpoint_read(path):
    """-> (sum [h, w, 3] f32, samples_done, seed)"""
    L = _bind_world(lib())
    w, h, n, seed, p = C.c_int32(), C.c_int32(), C.c_int32(), C.c_uint64(), C.POINTER(C.c_float)()
    _check(L.srt_checkpoint_read(os.fsencode(path), C.byref(w), C.byref(h), C.byref(n), C.byref(seed), C.byref(p)))
    a = np.ctypeslib.as_array(p, shape=(h.value * w.value * 3,)).copy().reshape(h.value, w.value, 3)
    L.srt_host_free(p)
    return a, n.value, seed.value
