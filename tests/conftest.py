import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from sexy_raytracer_b200.scene import Scene  # noqa: E402

# the tests push small scenes through the large-scene code paths with the library's tuning overrides (srt_launch.cuh: tune_env)
os.environ["SRT_DEBUG_TUNING"] = "1"

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = ["c1_shirley", "c2_cube", "c3_masterchief", "c3s_globe"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def scenes():
    return {n: Scene.load(os.path.join(ROOT, "scenes", n + ".npz")) for n in SCENES}


@pytest.fixture(scope="session")
def golden():
    class G:
        def kat(self, name):
            return np.load(os.path.join(GOLDEN, name + "_kat.npz"))

        def image(self, name):
            return np.load(os.path.join(GOLDEN, name + "_image.npz"))

        def aabb(self):
            return np.load(os.path.join(GOLDEN, "aabb_kat.npz"))

        def __call__(self, name):
            return np.load(os.path.join(GOLDEN, name + "_kat.npz"))
    return G()


@pytest.fixture(scope="session")
def ext_scene(scenes):
    """Config 3's scene with the two things the reference does not do switched on: per-vertex normals
    (the face normal bent towards the direction from the mesh's centroid to the vertex) and, for every textured PBR material,
    its base-colour image used as a packed metallicRoughnessTexture (SRT_MATF_PACKED_MR) with factors 0.9 / 0.8."""
    import copy
    from sexy_raytracer_b200 import abi
    s = copy.copy(scenes["c3_masterchief"])
    F = np.float32
    v = [s.tri_v0.astype(np.float64), s.tri_v1.astype(np.float64), s.tri_v2.astype(np.float64)]
    fn = np.cross(v[1] - v[0], v[2] - v[0])
    fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
    cen = np.concatenate(v).mean(axis=0)
    out = []
    for k in range(3):
        d = v[k] - cen
        d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30)
        n = fn + 0.6 * d
        out.append((n / np.linalg.norm(n, axis=1, keepdims=True)).astype(F))
    s.tri_n0, s.tri_n1, s.tri_n2 = out
    s.mat_tex = s.mat_tex.copy(); s.mat_params = s.mat_params.copy()
    s.mat_flags = np.zeros(len(s.mat_kind), np.int32)
    for i in range(len(s.mat_kind)):
        a = int(s.mat_tex[i, 0])
        if s.mat_kind[i] == abi.SRT_MAT_PBR and a >= 0 and s.tex_kind[a] == abi.SRT_TEX_IMAGE and s.tex_whb[a, 2] == 3:
            s.mat_tex[i, 2], s.mat_tex[i, 3] = a, -1
            s.mat_params[i, 0], s.mat_params[i, 1] = 0.9, 0.8
            s.mat_flags[i] = abi.MATF_PACKED_MR
    assert s.mat_flags.any()
    return s


@pytest.fixture(scope="session")
def tan_scene(ext_scene):
    """ext_scene plus per-vertex tangents (srt_scene_desc::tri_tan0..2, glTF TANGENT - not read by the reference): a tangent
    field perpendicular to the vertex normals; handedness +1 / -1 by triangle, and every third triangle without tangents
    (w = 0: it keeps calcTangentBasis' per-face basis)."""
    import copy
    s = copy.copy(ext_scene)
    F = np.float32
    out = []
    idx = np.arange(s.n_tri)
    w = np.where(idx % 3 == 0, 1.0, np.where(idx % 3 == 1, -1.0, 0.0)).astype(F)
    for n in (s.tri_n0, s.tri_n1, s.tri_n2):
        n = n.astype(np.float64)
        axis = np.where(np.abs(n[:, 1:2]) < 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
        t = np.cross(axis, n)
        t /= np.maximum(np.linalg.norm(t, axis=1, keepdims=True), 1e-30)
        t4 = np.concatenate([t, w[:, None]], axis=1).astype(F)
        t4[w == 0] = 0
        out.append(t4)
    s.tri_tan0, s.tri_tan1, s.tri_tan2 = out
    return s


@pytest.fixture(scope="session")
def port():
    """The plain-C restatement (checker).  Built on demand: gcc is part of the image."""
    from oracle import refbind
    if not refbind.PortOracle.available():
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libsrt_oracle.so"])
    return refbind.PortOracle()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (checker); only where oracle/_ref was built (needs /root/reference)."""
    from oracle import refbind
    if not refbind.RefOracle.available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (no /root/reference on this box)")
    return refbind.RefOracle()


@pytest.fixture(scope="session")
def device():
    from sexy_raytracer_b200.device import Device
    d = Device(0)
    yield d
    d.close()


@pytest.fixture(scope="session")
def dev_scenes(device, scenes):
    out = {n: device.upload(s) for n, s in scenes.items()}
    yield out
    for s in out.values():
        s.close()


def same_bits(a, b):
    """Bitwise equality of float arrays, NaN == NaN of the same payload class."""
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and (a.view(np.uint8) == b.view(np.uint8)).all()


def rel_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-30)


def psnr_u8(a, b, mask=None):
    a, b = a[..., :3].astype(np.float64), b[..., :3].astype(np.float64)
    d = (a - b) ** 2
    if mask is not None:
        d = d[mask]
    mse = d.mean()
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def tex_oob_mask(scene, tex, uvp):
    """Lookups where imagePNG::value reads px[1..2] past the end of a < 3 bytes-per-texel map
    (texture.h:145-147): undefined in the reference (heap bytes), pinned to 0 in the oracle and on the GPU."""
    w, h, bpp = (int(x) for x in scene.tex_whb[tex])
    if scene.tex_kind[tex] != 2 or bpp >= 3 or w == 0:
        return np.zeros(len(uvp), bool)
    F = np.float32
    u = np.clip(uvp[:, 0], 0, 1).astype(F)
    v = (F(1) - np.clip(uvp[:, 1], 0, 1).astype(F)).astype(F)
    i = np.minimum((u * F(w)).astype(np.int64), w - 1)
    j = np.minimum((v * F(h)).astype(np.int64), h - 1)
    return (j * w + i) * bpp + 2 >= w * h * bpp
