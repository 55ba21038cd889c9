"""Flattened scene container: the SoA buffers + 32-byte BVH nodes + material/texture tables that cross
the C ABI (include/srt.h: srt_scene_desc), held as numpy arrays, with .npz persistence.

A scene file also carries the camera (the ten derived members of the reference's camera.h:52-62, as
produced by the host-side camera class) and the render parameters of its config, so bench and tests
load one file per config.  No arithmetic on the path happens here.
"""
import ctypes as C
import numpy as np

from . import abi

_F = np.float32
_I = np.int32


def _f(a, cols):
    a = np.ascontiguousarray(a, dtype=_F)
    return a.reshape(-1, cols) if cols > 1 else a.reshape(-1)


class Scene:
    """Field names follow srt_scene_desc."""

    ARRAYS = ("tri_v0", "tri_v1", "tri_v2", "tri_uv0", "tri_uv1", "tri_uv2", "tri_mat",
              "sph_c0", "sph_c1", "sph_t0", "sph_t1", "sph_r", "sph_mat", "nodes",
              "mat_kind", "mat_tex", "mat_albedo", "mat_params",
              "tex_kind", "tex_color", "tex_evenodd", "tex_whb", "tex_missing")
    # not the reference's behaviour, absent from files that do not use them
    OPTIONAL = ("tri_n0", "tri_n1", "tri_n2", "tri_tan0", "tri_tan1", "tri_tan2", "mat_flags")

    def __init__(self):
        self.tri_v0 = self.tri_v1 = self.tri_v2 = np.zeros((0, 3), _F)
        self.tri_uv0 = self.tri_uv1 = self.tri_uv2 = np.zeros((0, 2), _F)
        self.tri_mat = np.zeros(0, _I)
        self.sph_c0 = self.sph_c1 = np.zeros((0, 3), _F)
        self.sph_t0 = self.sph_t1 = self.sph_r = np.zeros(0, _F)
        self.sph_mat = np.zeros(0, _I)
        self.nodes = np.zeros(0, abi.NODE_DTYPE)
        self.mat_kind = np.zeros(0, _I)
        self.mat_tex = np.zeros((0, 4), _I)       # albedo, normal, metallic, roughness texture ids
        self.mat_albedo = np.zeros((0, 4), _F)
        self.mat_params = np.zeros((0, 4), _F)    # metalness, roughness, fuzz, ir
        self.tex_kind = np.zeros(0, _I)
        self.tex_color = np.zeros((0, 3), _F)
        self.tex_evenodd = np.zeros((0, 2), _I)
        self.tex_whb = np.zeros((0, 3), _I)       # width, height, bytes per texel
        self.tex_missing = np.zeros(0, _I)        # 1 = image file was missing (data NULL)
        self.tex_data = {}                        # texture id -> uint8 array [h*w*bpp]
        self.tri_n0 = self.tri_n1 = self.tri_n2 = None   # per-vertex normals [n_tri,3], all three or none
        self.tri_tan0 = self.tri_tan1 = self.tri_tan2 = None   # per-vertex tangents xyz + handedness [n_tri,4], all three or none
        self.mat_flags = None                     # [n_mat] abi.MATF_*, None = all 0
        # config
        self.camera = np.zeros(24, _F)            # srt_camera, flat
        self.cam_args = np.zeros(15, _F)          # eye3 lookAt3 up3 vfov aspect aperture focus t0 t1
        self.render = np.zeros(4, _I)             # width height spp max_bounce
        self.background = np.array([0.53, 0.81, 0.92], _F)   # main.cpp:170
        self.name = ""
        self._keep = []

    # ---- sizes
    @property
    def n_tri(self):
        return int(self.tri_mat.shape[0])

    @property
    def n_sph(self):
        return int(self.sph_mat.shape[0])

    @property
    def n_prim(self):
        return self.n_tri + self.n_sph

    def host_bytes(self):
        """Bytes handed across the ABI by srt_scene_create (the H2D volume of one scene upload)."""
        n = sum(getattr(self, k).nbytes for k in self.ARRAYS)
        n += sum(getattr(self, k).nbytes for k in self.OPTIONAL if getattr(self, k) is not None)
        return int(n + sum(v.nbytes for v in self.tex_data.values()))

    # ---- C view
    def desc(self):
        """-> abi.SceneDesc whose pointers stay valid while this Scene is alive and unmodified."""
        keep = []

        def fp(a):
            a = np.ascontiguousarray(a, dtype=_F); keep.append(a)
            return a.ctypes.data_as(abi.c_float_p)

        def ip(a):
            a = np.ascontiguousarray(a, dtype=_I); keep.append(a)
            return a.ctypes.data_as(abi.c_int32_p)

        d = abi.SceneDesc()
        d.n_tri = self.n_tri
        d.tri_v0, d.tri_v1, d.tri_v2 = fp(self.tri_v0), fp(self.tri_v1), fp(self.tri_v2)
        d.tri_uv0, d.tri_uv1, d.tri_uv2 = fp(self.tri_uv0), fp(self.tri_uv1), fp(self.tri_uv2)
        d.tri_mat = ip(self.tri_mat)
        if self.tri_n0 is not None:
            assert all(np.shape(x) == (self.n_tri, 3) for x in (self.tri_n0, self.tri_n1, self.tri_n2))
            d.tri_n0, d.tri_n1, d.tri_n2 = fp(self.tri_n0), fp(self.tri_n1), fp(self.tri_n2)
        if self.tri_tan0 is not None:
            assert all(np.shape(x) == (self.n_tri, 4) for x in (self.tri_tan0, self.tri_tan1, self.tri_tan2))
            d.tri_tan0, d.tri_tan1, d.tri_tan2 = fp(self.tri_tan0), fp(self.tri_tan1), fp(self.tri_tan2)
        d.n_sph = self.n_sph
        d.sph_c0, d.sph_c1 = fp(self.sph_c0), fp(self.sph_c1)
        d.sph_t0, d.sph_t1, d.sph_r = fp(self.sph_t0), fp(self.sph_t1), fp(self.sph_r)
        d.sph_mat = ip(self.sph_mat)
        nodes = np.ascontiguousarray(self.nodes, dtype=abi.NODE_DTYPE); keep.append(nodes)
        d.n_nodes = nodes.shape[0]
        d.nodes = nodes.ctypes.data_as(C.POINTER(abi.BvhNode))
        nm = self.mat_kind.shape[0]
        mats = (abi.Material * max(nm, 1))()
        for i in range(nm):
            m = mats[i]
            m.kind = int(self.mat_kind[i])
            m.albedo_tex, m.normal_tex, m.metallic_tex, m.roughness_tex = (int(x) for x in self.mat_tex[i])
            m.albedo[:] = [float(x) for x in self.mat_albedo[i]]
            m.metalness, m.roughness, m.fuzz, m.ir = (float(x) for x in self.mat_params[i])
            m.flags = int(self.mat_flags[i]) if self.mat_flags is not None else 0
        d.n_mat, d.materials = nm, mats
        nt = self.tex_kind.shape[0]
        texs = (abi.Texture * max(nt, 1))()
        for i in range(nt):
            t = texs[i]
            t.kind = int(self.tex_kind[i])
            t.color[:] = [float(x) for x in self.tex_color[i]]
            t.even, t.odd = int(self.tex_evenodd[i][0]), int(self.tex_evenodd[i][1])
            t.width, t.height, t.bpp = (int(x) for x in self.tex_whb[i])
            if t.kind == abi.SRT_TEX_IMAGE and not self.tex_missing[i]:
                a = np.ascontiguousarray(self.tex_data[i], dtype=np.uint8); keep.append(a)
                assert a.size == t.width * t.height * t.bpp
                t.data = a.ctypes.data_as(abi.c_uint8_p)
        d.n_tex, d.textures = nt, texs
        keep += [mats, texs]
        self._keep = keep
        return d

    def camera_struct(self):
        cam = abi.Camera()
        C.memmove(C.byref(cam), np.ascontiguousarray(self.camera, _F).ctypes.data, C.sizeof(cam))
        return cam

    def render_params(self, **over):
        p = abi.RenderParams()
        p.width, p.height, p.spp, p.max_bounce = (int(x) for x in self.render)
        p.background[:] = [float(x) for x in self.background]
        p.sample_begin, p.sample_count, p.seed, p.rr_begin, p.variant, p.sampler = 0, 0, 1, 0, 0, 0
        for k, v in over.items():
            setattr(p, k, v)
        return p

    # ---- persistence
    def save(self, path):
        d = {k: getattr(self, k) for k in self.ARRAYS}
        d.update({k: getattr(self, k) for k in self.OPTIONAL if getattr(self, k) is not None})
        # identical images (the same PNG loaded by several materials) are stored once
        blobs, ref = [], np.full(self.tex_kind.shape[0], -1, _I)
        for i, a in sorted(self.tex_data.items()):
            a = np.ascontiguousarray(a, np.uint8)
            for k, b in enumerate(blobs):
                if b.shape == a.shape and np.array_equal(a, b):
                    ref[i] = k
                    break
            else:
                ref[i] = len(blobs)
                blobs.append(a)
        for k, b in enumerate(blobs):
            d["tex_blob_%d" % k] = b
        d.update(tex_blob_ref=ref, camera=self.camera, cam_args=self.cam_args, render=self.render,
                 background=self.background, name=np.array(self.name))
        np.savez_compressed(path, **d)

    @classmethod
    def load(cls, path):
        s = cls()
        with np.load(path, allow_pickle=False) as z:
            for k in cls.ARRAYS:
                setattr(s, k, z[k])
            for k in cls.OPTIONAL:
                if k in z.files:
                    setattr(s, k, z[k])
            ref = z["tex_blob_ref"]
            blobs = {}
            for i, k in enumerate(ref):
                if k >= 0:
                    if int(k) not in blobs:
                        blobs[int(k)] = z["tex_blob_%d" % k]
                    s.tex_data[i] = blobs[int(k)]
            s.camera, s.cam_args, s.render, s.background = z["camera"], z["cam_args"], z["render"], z["background"]
            s.name = str(z["name"])
        return s
