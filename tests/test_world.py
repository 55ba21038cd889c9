"""The host side of the drop-in (include/srt_world.hpp -> libsrt_host.so): the scene classes with the reference's
constructor signatures, the glTF reader, the PNG codec, the flattening.  CPU tests; the last one drives the example
main.cpp-equivalent on the GPU."""
import base64
import ctypes as C
import io
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SCENES
from sexy_raytracer_b200 import abi, host
from sexy_raytracer_b200.scene import Scene

REF_DATA = "/root/reference/data"
needs_ref_data = pytest.mark.skipif(not os.path.isdir(REF_DATA), reason="the reference's data/ directory is not on this box")


def _same(a, b):
    return a.shape == b.shape and a.tobytes() == b.tobytes()


# ---- the four reference scenes, built by the C++ host classes, equal the compiled reference's own export -------------
@needs_ref_data
@pytest.mark.parametrize("which,name", [(1, "c1_shirley"), (2, "c2_cube"), (3, "c3_masterchief"), (4, "c3s_globe")])
def test_host_classes_reproduce_reference_flattening(which, name, scenes):
    """scenes/*.npz were exported by the compiled reference (oracle/make_golden.py -> ref_scene_export).  The same scene
    written against srt_world.hpp (host/srt_scenes.cpp mirrors main.cpp:82-146 statement for statement) must flatten to
    the identical bytes: geometry read by the glTF reader, texel bytes decoded by the PNG reader, the random sphere
    parameters of main.cpp:92-122 (same generator, same draw order), material / texture tables, and the tree."""
    s, draws = host.reference_scene(which, REF_DATA)
    g = scenes[name]
    for k in Scene.ARRAYS:
        assert _same(getattr(s, k), getattr(g, k)), k
    assert set(s.tex_data) == set(g.tex_data)
    for i in g.tex_data:
        assert np.array_equal(s.tex_data[i].ravel(), np.asarray(g.tex_data[i]).ravel()), "texel bytes of texture %d" % i
    # one draw per node for the tree + the draws of the scene block (main.cpp:92-122 for config 1)
    assert draws >= len(g.nodes)
    if which != 1:
        assert draws == len(g.nodes)


def test_reference_scene_errors_are_loud(tmp_path):
    with pytest.raises(RuntimeError):
        host.reference_scene(3, str(tmp_path))          # no glTF there
    with pytest.raises(RuntimeError):
        host.reference_scene(9, str(tmp_path))
    s, draws = host.reference_scene(1, str(tmp_path))   # config 1 needs no files
    assert s.n_sph == 485 and s.n_tri == 0 and len(s.nodes) == 511


# ---- PNG ---------------------------------------------------------------------------------------------------------------
def _pil():
    return pytest.importorskip("PIL.Image")


def _stb(ref_so):
    L = C.CDLL(ref_so)
    L.stbi_load_from_memory.restype = C.POINTER(C.c_uint8)
    L.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    L.stbi_image_free.argtypes = [C.c_void_p]

    def load(data, req):
        w, h, n = C.c_int(), C.c_int(), C.c_int()
        p = L.stbi_load_from_memory(data, len(data), C.byref(w), C.byref(h), C.byref(n), req)
        assert p, "stb could not decode"
        a = np.ctypeslib.as_array(p, shape=(h.value, w.value, req)).copy()
        L.stbi_image_free(p)
        return a
    return load


def _png_cases():
    Image = _pil()
    rng = np.random.default_rng(5)
    out = {}
    rgb = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    rgba = rng.integers(0, 256, (19, 31, 4), dtype=np.uint8)
    grey = rng.integers(0, 256, (23, 29), dtype=np.uint8)
    smooth = (np.add.outer(np.arange(64), np.arange(80)) * 2 % 256).astype(np.uint8)       # exercises every row filter
    for name, im, kw in (("rgb", Image.fromarray(rgb, "RGB"), {}), ("rgba", Image.fromarray(rgba, "RGBA"), {}),
                         ("grey", Image.fromarray(grey, "L"), {}), ("smooth", Image.fromarray(np.dstack([smooth] * 3), "RGB"), {"optimize": True}),
                         ("la", Image.fromarray(np.dstack([grey, grey[::-1]]), "LA"), {}),
                         ("pal", Image.fromarray(rgb, "RGB").quantize(16), {}), ("pal256", Image.fromarray(rgb, "RGB").quantize(200), {}),
                         ("bw", Image.fromarray(grey > 127), {}),
                         ("g16", Image.fromarray((grey.astype(np.uint16) * 257 + 13).astype(np.uint16)), {})):
        b = io.BytesIO()
        im.save(b, "PNG", **kw)
        out[name] = b.getvalue()
    return out


def test_png_read_matches_pil():
    Image = _pil()
    for name, data in _png_cases().items():
        im = Image.open(io.BytesIO(data))
        if name in ("rgb", "smooth", "pal", "pal256"):
            want = np.asarray(im.convert("RGB"))
            assert np.array_equal(host.png_read(data, 3), want), name
        if name in ("grey", "bw"):
            want = np.asarray(im.convert("L"))
            assert np.array_equal(host.png_read(data, 1)[:, :, 0], want), name
            assert np.array_equal(host.png_read(data, 3), np.dstack([want] * 3)), name
        if name == "rgba":
            assert np.array_equal(host.png_read(data, 4), np.asarray(im)), name
            assert np.array_equal(host.png_read(data, 3), np.asarray(im)[:, :, :3]), name


def test_png_read_matches_stb(ref):
    """Channel conversion and bit-depth rules are stb_image's (what imagePNG gets from stbi_load, texture.h:115): checked
    against the stb_image compiled into the reference checker, for every colour type x requested channel count."""
    from oracle import refbind
    load = _stb(refbind.REF_SO)
    for name, data in _png_cases().items():
        for req in (1, 2, 3, 4):
            assert np.array_equal(host.png_read(data, req), load(data, req)), (name, req)


@needs_ref_data
def test_png_read_reference_assets_vs_stb(ref):
    from oracle import refbind
    load = _stb(refbind.REF_SO)
    for f, req in (("Image_0.png", 3), ("Image_1.png", 3), ("worldisyoursbanner.png", 1), ("rustediron2_metallic.png", 1),
                   ("rustediron2_roughness.png", 3)):
        data = open(os.path.join(REF_DATA, f), "rb").read()
        assert np.array_equal(host.png_read(os.path.join(REF_DATA, f), req), load(data, req)), f


def test_png_write_round_trip(tmp_path):
    Image = _pil()
    rng = np.random.default_rng(1)
    for comp in (1, 2, 3, 4):
        a = rng.integers(0, 256, (45, 61, comp), dtype=np.uint8)
        p = str(tmp_path / ("w%d.png" % comp))
        host.png_write(p, a)
        assert np.array_equal(host.png_read(p, comp), a)
        b = np.asarray(Image.open(p))                      # an independent decoder reads the file too
        assert np.array_equal(b.reshape(a.shape), a)
    with pytest.raises(RuntimeError):
        host.png_read(str(tmp_path / "missing.png"), 3)
    bad = tmp_path / "bad.png"
    bad.write_bytes(b"\x89PNG\r\n\x1a\n" + b"junk" * 10)
    with pytest.raises(RuntimeError):
        host.png_read(str(bad), 3)


# ---- glTF ----------------------------------------------------------------------------------------------------------------
def _tiny_gltf(tmp_path, glb=False, idx_type=5123, acc_offset=0, translate=None, name="t"):
    """Two triangles (a unit quad in z = 0) with UVs; optional accessor byteOffset, 32-bit indices, node translation."""
    pos = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32)
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    idx = np.array([0, 1, 2, 0, 2, 3], {5123: np.uint16, 5125: np.uint32, 5121: np.uint8}[idx_type])
    pad = b"\xee" * acc_offset
    blob = pad + pos.tobytes()
    v0 = (0, len(blob))
    o1 = len(blob); blob += uv.tobytes(); v1 = (o1, len(blob) - o1)
    o2 = len(blob); blob += idx.tobytes(); v2 = (o2, len(blob) - o2)
    blob += b"\0" * (-len(blob) % 4)
    node = {"mesh": 0}
    if translate is not None:
        node["translation"] = list(translate)
    doc = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}], "nodes": [node],
           "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.25, 0.5, 0.75, 1.0], "metallicFactor": 0.5,
                                                   "roughnessFactor": 0.125}}],
           "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "TEXCOORD_0": 1}, "indices": 2, "material": 0}]}],
           "accessors": [{"bufferView": 0, "byteOffset": acc_offset, "componentType": 5126, "count": 4, "type": "VEC3"},
                         {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC2"},
                         {"bufferView": 2, "componentType": idx_type, "count": 6, "type": "SCALAR"}],
           "bufferViews": [{"buffer": 0, "byteOffset": v0[0], "byteLength": v0[1]}, {"buffer": 0, "byteOffset": v1[0], "byteLength": v1[1]},
                           {"buffer": 0, "byteOffset": v2[0], "byteLength": v2[1]}],
           "buffers": [{"byteLength": len(blob)}]}
    if glb:
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(blob), 0x004E4942) + blob
        p = tmp_path / (name + ".glb")
        p.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        doc["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(blob).decode()
        p = tmp_path / (name + ".gltf")
        p.write_text(json.dumps(doc, indent=1))
    return str(p), pos, uv


@pytest.mark.parametrize("glb", [False, True])
def test_gltf_reader_basic(tmp_path, glb):
    path, pos, uv = _tiny_gltf(tmp_path, glb=glb)
    s = host.gltf_scene(path)
    assert s.n_tri == 2 and s.n_sph == 0 and len(s.nodes) == 1
    assert np.array_equal(s.tri_v0, pos[[0, 0]]) and np.array_equal(s.tri_v1, pos[[1, 2]]) and np.array_equal(s.tri_v2, pos[[2, 3]])
    assert np.array_equal(s.tri_uv1, uv[[1, 2]])
    assert s.mat_kind.tolist() == [abi.SRT_MAT_PBR] and s.mat_tex.tolist() == [[-1, -1, -1, -1]]
    assert np.allclose(s.mat_albedo[0], [0.25, 0.5, 0.75, 1.0]) and s.mat_params[0, 0] == 0.5 and s.mat_params[0, 1] == 0.125
    assert sorted((~s.nodes["left"][0], ~s.nodes["right"][0])) == [0, 1]


def test_gltf_reader_quirks_and_fixes(tmp_path):
    """the reference reads from bufferView.byteOffset only, 16-bit indices only, no node transforms
    (model.h:335-365,444-452).  referenceQuirks=True reproduces that; False honours accessor.byteOffset, the index
    component type and node TRS."""
    path, pos, _ = _tiny_gltf(tmp_path, acc_offset=12, idx_type=5125, translate=(10, 20, 30), name="q")
    fixed = host.gltf_scene(path, reference_quirks=False)
    assert fixed.n_tri == 2
    assert np.array_equal(fixed.tri_v0, (pos + np.float32([10, 20, 30]))[[0, 0]])
    assert np.array_equal(fixed.tri_v2, (pos + np.float32([10, 20, 30]))[[2, 3]])
    quirky = host.gltf_scene(path, reference_quirks=True)
    # 32-bit indices read as 16-bit: (0,0,1), (0,2,0) -> two degenerate triangles; positions start 12 bytes early
    assert quirky.n_tri == 2 and not np.array_equal(quirky.tri_v0, fixed.tri_v0)
    with pytest.raises(RuntimeError):
        host.gltf_scene(str(tmp_path / "nope.gltf"))
    bad = tmp_path / "bad.gltf"
    bad.write_text("{ not json")
    with pytest.raises(RuntimeError):
        host.gltf_scene(str(bad))


def test_gltf_smooth_normals_and_metallic_roughness_options(tmp_path):
    """SURVEY 8.6 next-3, the two things the reference never does: NORMAL is read (and carried through node transforms by the
    inverse transpose) only under smoothNormals; a metallicRoughnessTexture is sampled only under sampleMetallicRoughness.
    With both off the flattened scene is what the reference's loader would give."""
    path, pos, _ = _tiny_gltf(tmp_path, name="n")
    doc = json.loads(open(path).read())
    nrm = np.array([[0, 0, 1], [0.6, 0, 0.8], [0, 0.6, 0.8], [-0.6, 0, 0.8]], np.float32)
    blob = base64.b64decode(doc["buffers"][0]["uri"].split(",", 1)[1])
    off = len(blob)
    blob += nrm.tobytes()
    doc["buffers"][0] = {"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}
    doc["bufferViews"].append({"buffer": 0, "byteOffset": off, "byteLength": nrm.nbytes})
    doc["accessors"].append({"bufferView": 3, "componentType": 5126, "count": 4, "type": "VEC3"})
    doc["meshes"][0]["primitives"][0]["attributes"]["NORMAL"] = 3
    doc["nodes"][0]["scale"] = [2.0, 1.0, 1.0]
    mr = np.zeros((2, 2, 3), np.uint8)
    mr[..., 1], mr[..., 2] = [[10, 20], [30, 40]], [[200, 210], [220, 230]]
    host.png_write(str(tmp_path / "mr.png"), mr)
    doc["images"] = [{"uri": "mr.png"}]
    doc["textures"] = [{"source": 0}]
    doc["materials"][0]["pbrMetallicRoughness"]["metallicRoughnessTexture"] = {"index": 0}
    open(path, "w").write(json.dumps(doc))
    tex_dir = str(tmp_path) + "/"
    plain = host.gltf_scene(path, tex_dir, reference_quirks=False)
    assert plain.tri_n0 is None and plain.mat_flags is None and plain.mat_tex[0].tolist() == [-1, -1, -1, -1]
    s = host.gltf_scene(path, tex_dir, reference_quirks=False, smooth_normals=True, sample_metallic_roughness=True)
    assert _same(s.tri_v0, plain.tri_v0) and _same(s.nodes, plain.nodes)
    want = nrm.astype(np.float64) / np.array([2.0, 1.0, 1.0])          # inverse transpose of diag(2,1,1)
    want = (want / np.linalg.norm(want, axis=1, keepdims=True)).astype(np.float32)
    assert np.allclose(s.tri_n0, want[[0, 0]], atol=1e-6) and np.allclose(s.tri_n1, want[[1, 2]], atol=1e-6)
    assert np.allclose(s.tri_n2, want[[2, 3]], atol=1e-6)
    assert s.mat_flags.tolist() == [abi.MATF_PACKED_MR]
    t = int(s.mat_tex[0, 2])
    assert t >= 0 and s.mat_tex[0, 3] == -1 and s.tex_whb[t].tolist() == [2, 2, 3]
    assert np.array_equal(np.asarray(s.tex_data[t]).reshape(2, 2, 3), mr)
    assert s.mat_params[0, 0] == 0.5 and s.mat_params[0, 1] == 0.125     # the factors still multiply
    # under the reference's quirks the normals are read the same way, without the node transform
    q = host.gltf_scene(path, tex_dir, reference_quirks=True, smooth_normals=True)
    assert np.allclose(q.tri_n1, nrm[[1, 2]], atol=1e-7) and q.mat_flags is None


@needs_ref_data
def test_gltf_reference_assets_both_modes():
    """The reference's own assets carry no accessor offsets / strides / node transforms that matter, so both modes agree."""
    for f in ("cube.gltf", "square.gltf", "masterchief2-separate-xf.gltf"):
        a = host.gltf_scene(os.path.join(REF_DATA, f), REF_DATA + "/", True)
        b = host.gltf_scene(os.path.join(REF_DATA, f), REF_DATA + "/", False)
        for k in ("tri_v0", "tri_v1", "tri_v2", "tri_uv0", "tri_uv1", "tri_uv2", "tri_mat"):
            assert _same(getattr(a, k), getattr(b, k)), (f, k)


# ---- ordered builder -----------------------------------------------------------------------------------------------------
def test_bvh_build_ordered(scenes):
    s = scenes["c2_cube"]
    L = host._bind_world(host.lib())
    d = s.desc()
    n = s.n_prim

    def build(order):
        out, cnt, draws = C.POINTER(abi.BvhNode)(), C.c_int64(), C.c_uint64()
        o = None if order is None else np.ascontiguousarray(order, np.int32)
        rc = L.srt_bvh_build_ordered(C.byref(d), None if o is None else o.ctypes.data_as(C.c_void_p), 0.0, 1.0, 0, 1,
                                     C.byref(out), C.byref(cnt), C.byref(draws))
        if rc != 0:
            raise RuntimeError(L.srt_host_last_error().decode())
        a = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_uint8)), shape=(cnt.value * 32,)).view(abi.NODE_DTYPE).copy()
        L.srt_bvh_free(out)
        return a

    ident = build(np.arange(n))
    assert _same(ident, build(None)) and _same(ident, s.nodes)
    rev = build(np.arange(n)[::-1])
    leaves = np.concatenate([~rev["left"][rev["left"] < 0], ~rev["right"][(rev["right"] < 0) & (rev["right"] != abi.SRT_BVH_NONE)]])
    assert sorted(leaves.tolist()) == list(range(n))               # every primitive exactly once
    assert _same(rev["bmin"][0], ident["bmin"][0]) and _same(rev["bmax"][0], ident["bmax"][0])
    with pytest.raises(RuntimeError):
        build(np.zeros(n))                                           # not a permutation


# ---- the example program: the reference's main.cpp over the host classes + the device ---------------------------------
def test_example_builds():
    from sexy_raytracer_b200 import build
    build.build_host()
    exe = build.build_example()
    assert os.access(exe, os.X_OK)
    # without a CUDA library / GPU it must fail loudly, not fall back
    r = subprocess.run([exe, "shirley", "45", "1", os.devnull], capture_output=True, text=True,
                       env=dict(os.environ, SRT_B200_LIBRARY="/nonexistent/libsrt_b200.so"))
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_example_renders_config1_like_the_library(tmp_path, device, dev_scenes, scenes):
    """examples/sexy-raytracer shirley: scene built by the C++ classes, rendered through b200Device (dlopen of
    libsrt_b200.so), written by srt_png_write == the frame srt_render gives for scenes/c1_shirley.npz (same seed)."""
    from sexy_raytracer_b200 import build
    exe = build.build_example()
    out = str(tmp_path / "test.png")
    r = subprocess.run([exe, "shirley", "225", "16", out], capture_output=True, text=True,
                       env=dict(os.environ, SRT_B200_LIBRARY=build.LIB))
    assert r.returncode == 0, r.stderr
    got = host.png_read(out, 4)
    s = scenes["c1_shirley"]
    assert got.shape == (225, 400, 4)
    _, want, _ = dev_scenes["c1_shirley"].render(s.render_params(width=400, height=225, spp=16, max_bounce=10, seed=1))
    want = np.asarray(want).reshape(225, 400, 4)
    assert (got[..., 3] == 255).all()
    # the same kernels on the same description and seed: identical up to the order of the atomic adds of the accumulation
    assert (np.abs(got.astype(int) - want.astype(int)) <= 1).mean() > 0.999


# ---- progressive accumulation + checkpoint ----------------------------------------------------------
def test_gltf_vertex_tangents_option(tmp_path):
    """glTF TANGENT (VEC4: xyz + handedness) is read only under vertexTangents - the reference never reads it
    (model.h:335-365) - carried through node transforms as a direction (upper 3x3, renormalised; w untouched), and lands in
    srt_scene_desc::tri_tan0..2 per triangle vertex.  With the flag off the flattened scene is unchanged."""
    path, pos, _ = _tiny_gltf(tmp_path, name="t")
    doc = json.loads(open(path).read())
    tan = np.array([[1, 0, 0, 1], [0.6, 0.8, 0, 1], [0, 1, 0, -1], [0, 0.6, 0.8, -1]], np.float32)
    blob = base64.b64decode(doc["buffers"][0]["uri"].split(",", 1)[1])
    off = len(blob)
    blob += tan.tobytes()
    doc["buffers"][0] = {"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}
    doc["bufferViews"].append({"buffer": 0, "byteOffset": off, "byteLength": tan.nbytes})
    doc["accessors"].append({"bufferView": 3, "componentType": 5126, "count": 4, "type": "VEC4"})
    doc["meshes"][0]["primitives"][0]["attributes"]["TANGENT"] = 3
    doc["nodes"][0]["scale"] = [2.0, 1.0, 1.0]
    open(path, "w").write(json.dumps(doc))
    tex_dir = str(tmp_path) + "/"
    plain = host.gltf_scene(path, tex_dir, reference_quirks=False)
    assert plain.tri_tan0 is None
    s = host.gltf_scene(path, tex_dir, reference_quirks=False, vertex_tangents=True)
    assert _same(s.tri_v0, plain.tri_v0) and _same(s.nodes, plain.nodes) and s.tri_n0 is None
    want = tan.astype(np.float64).copy()
    want[:, 0] *= 2.0                                                    # a direction under diag(2,1,1)
    want[:, :3] /= np.linalg.norm(want[:, :3], axis=1, keepdims=True)
    want = want.astype(np.float32)
    assert np.allclose(s.tri_tan0, want[[0, 0]], atol=1e-6) and np.allclose(s.tri_tan1, want[[1, 2]], atol=1e-6)
    assert np.allclose(s.tri_tan2, want[[2, 3]], atol=1e-6)
    q = host.gltf_scene(path, tex_dir, reference_quirks=True, vertex_tangents=True)        # the quirks ignore node transforms
    assert np.allclose(q.tri_tan1, tan[[1, 2]], atol=1e-7)


def test_pfm_round_trip(tmp_path):
    """srt_pfm_write / srt_pfm_read (a raw fp32 dump for parity tooling): values bit-exact, row 0 stays the
    top row although the file stores rows bottom-up, the scale factor applies on the way out, the header is what other PFM
    readers expect."""
    rng = np.random.default_rng(3)
    a = rng.standard_normal((7, 5, 3)).astype(np.float32) * 100
    a[0, 0] = [np.inf, -0.0, 1e-38]
    f = str(tmp_path / "x.pfm")
    host.pfm_write(f, a)
    raw = open(f, "rb").read()
    assert raw.startswith(b"PF\n5 7\n-1.0\n")
    body = np.frombuffer(raw[len(b"PF\n5 7\n-1.0\n"):], "<f4").reshape(7, 5, 3)
    assert (body[::-1].view(np.uint32) == a.view(np.uint32)).all()          # bottom row first in the file
    b = host.pfm_read(f)
    assert b.shape == a.shape and (b.view(np.uint32) == a.view(np.uint32)).all()
    host.pfm_write(f, a[1:], scale=0.25)
    assert np.array_equal(host.pfm_read(f), a[1:] * np.float32(0.25))
    open(f, "wb").write(b"P6\n5 7\n255\n")
    with pytest.raises(Exception):
        host.pfm_read(f)


def test_checkpoint_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    a = rng.random((7, 11, 3), dtype=np.float32) * 100
    p = str(tmp_path / "film.srtf")
    host.checkpoint_write(p, a, 123, seed=0xDEADBEEF12345)
    b, n, seed = host.checkpoint_read(p)
    assert n == 123 and seed == 0xDEADBEEF12345 and b.tobytes() == a.tobytes()
    (tmp_path / "short.srtf").write_bytes(open(p, "rb").read()[:-8])
    with pytest.raises(RuntimeError):
        host.checkpoint_read(str(tmp_path / "short.srtf"))
    (tmp_path / "junk.srtf").write_bytes(b"PNG?" * 20)
    with pytest.raises(RuntimeError):
        host.checkpoint_read(str(tmp_path / "junk.srtf"))


@pytest.mark.gpu
def test_film_progressive_equals_one_call(device, dev_scenes, scenes, tmp_path):
    """K calls of n samples = one call of K*n samples (Philox is keyed by the global sample index), also across a
    checkpoint written by srt_checkpoint_write and restored into a new film."""
    from sexy_raytracer_b200.device import Film
    s, ds = scenes["c3_masterchief"], dev_scenes["c3_masterchief"]
    W, H = 160, 90
    want, want_rgba, st = ds.render(s.render_params(width=W, height=H, spp=16, seed=5))
    P = s.render_params(width=W, height=H, seed=5)
    film = Film(device, W, H)
    segs = 0
    for _ in range(2):
        segs += int(film.add_samples(ds, 4, P).segments)
    assert film.samples == 8
    half, _ = film.resolve(want_rgba=False)
    host.checkpoint_write(str(tmp_path / "f.srtf"), half, film.samples, seed=5)
    film.close()
    saved, done, seed = host.checkpoint_read(str(tmp_path / "f.srtf"))
    film2 = Film(device, W, H)
    film2.load(saved, done)
    for _ in range(2):
        segs += int(film2.add_samples(ds, 4, P).segments)
    assert film2.samples == 16 and segs == int(st.segments)
    got, got_rgba = film2.resolve()
    film2.close()
    fin = np.isfinite(want).all(axis=2)
    assert (np.isfinite(got).all(axis=2) == fin).all()
    assert np.abs(got[fin] - want[fin]).max() <= 2e-5 * max(1.0, np.abs(want[fin]).max())
    assert (np.abs(got_rgba.astype(int) - want_rgba.astype(int)) <= 1).all()
    with pytest.raises(Exception):
        Film(device, 0, 10)


@pytest.mark.gpu
def test_example_progressive_resumes(tmp_path):
    """examples/sexy-raytracer with passes: 16 samples in 4 passes, killed-and-resumed from the checkpoint = one 16-sample run."""
    from sexy_raytracer_b200 import build
    exe = build.build_example()
    env = dict(os.environ, SRT_B200_LIBRARY=build.LIB)
    one = str(tmp_path / "one.png")
    assert subprocess.run([exe, "shirley", "90", "16", one], capture_output=True, env=env).returncode == 0
    prog = str(tmp_path / "prog.png")
    assert subprocess.run([exe, "shirley", "90", "8", prog, "2"], capture_output=True, env=env).returncode == 0     # stops at 8
    _, done, _ = host.checkpoint_read(prog + ".srtf")
    assert done == 8
    r = subprocess.run([exe, "shirley", "90", "16", prog, "4"], capture_output=True, text=True, env=env)              # resumes to 16
    assert r.returncode == 0 and "resuming" in r.stderr
    _, done, _ = host.checkpoint_read(prog + ".srtf")
    assert done == 16
    a, b = host.png_read(one, 4), host.png_read(prog, 4)
    assert a.shape == b.shape and (np.abs(a.astype(int) - b.astype(int)) <= 1).mean() > 0.999


@pytest.mark.gpu
def test_example_refuses_to_mix_sample_streams(tmp_path):
    """a checkpoint records seed, sampler and depth (SRTF version 2): resuming it with another seed would continue the global
    sample indices with different draws - b200Device::rtFrameProgressive refuses."""
    from sexy_raytracer_b200 import build
    exe = build.build_example()
    env = dict(os.environ, SRT_B200_LIBRARY=build.LIB)
    prog = str(tmp_path / "mix.png")
    assert subprocess.run([exe, "shirley", "90", "8", prog, "2", "1"], capture_output=True, env=env).returncode == 0
    r = subprocess.run([exe, "shirley", "90", "16", prog, "4", "2"], capture_output=True, text=True, env=env)       # other seed
    assert r.returncode != 0 and "another seed" in r.stderr
    _, done, seed = host.checkpoint_read(prog + ".srtf")
    assert done == 8 and seed == 1                                 # the checkpoint is untouched


def test_checkpoint_version2_round_trip(tmp_path):
    import ctypes as C
    L = host._bind_world(host.lib())
    L.srt_checkpoint_write2.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.c_int32, C.c_int32, C.c_void_p]
    L.srt_checkpoint_read2.argtypes = [C.c_char_p] + [C.POINTER(C.c_int32)] * 3 + [C.POINTER(C.c_uint64)] + [C.POINTER(C.c_int32)] * 2 + \
        [C.POINTER(C.POINTER(C.c_float))]
    a = np.random.default_rng(1).random((5, 7, 3)).astype(np.float32)
    p = str(tmp_path / "v2.srtf").encode()
    assert L.srt_checkpoint_write2(p, 7, 5, 40, 99, 1, 6, a.ctypes.data_as(C.c_void_p)) == 0
    w, h, n, smp, mb = (C.c_int32() for _ in range(5))
    seed = C.c_uint64()
    out = C.POINTER(C.c_float)()
    assert L.srt_checkpoint_read2(p, C.byref(w), C.byref(h), C.byref(n), C.byref(seed), C.byref(smp), C.byref(mb), C.byref(out)) == 0
    assert (w.value, h.value, n.value, seed.value, smp.value, mb.value) == (7, 5, 40, 99, 1, 6)
    assert (np.ctypeslib.as_array(out, shape=(105,)) == a.ravel()).all()
    L.srt_host_free(out)
    b, n1, seed1 = host.checkpoint_read(p.decode())               # the version-1 entry point reads version-2 files too
    assert n1 == 40 and seed1 == 99 and (b == a).all()
    host.checkpoint_write(p.decode(), a, 3, seed=5)                # ... and what it writes carries "not recorded"
    assert L.srt_checkpoint_read2(p, C.byref(w), C.byref(h), C.byref(n), C.byref(seed), C.byref(smp), C.byref(mb), C.byref(out)) == 0
    assert (smp.value, mb.value) == (-1, -1)
    L.srt_host_free(out)


def test_gltf_rejects_malformed_accessors(tmp_path):
    """accessor / bufferView numbers come from an untrusted file: negative, fractional, huge and wrapping values, strides below
    the element size and views past the buffer are rejected, not read."""
    import base64
    import json
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32).tobytes()
    idx = np.array([0, 1, 2], np.uint16).tobytes()
    blob = pos + idx + b"\0\0"

    def doc(**over):
        d = {"asset": {"version": "2.0"}, "buffers": [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}],
             "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 36}, {"buffer": 0, "byteOffset": 36, "byteLength": 6}],
             "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"},
                           {"bufferView": 1, "componentType": 5123, "count": 3, "type": "SCALAR"}],
             "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [1, 1, 1, 1]}}],
             "meshes": [{"primitives": [{"attributes": {"POSITION": 0}, "indices": 1, "material": 0}]}], "nodes": [{"mesh": 0}], "scenes": [{"nodes": [0]}]}
        for path, v in over.items():
            a, i, k = path.split("/")
            d[a][int(i)][k] = v
        return d

    def load(d, quirks):
        p = str(tmp_path / "m.gltf")
        open(p, "w").write(json.dumps(d))
        try:
            return host.gltf_scene(p, str(tmp_path) + "/", reference_quirks=quirks)
        except Exception:
            return None
    good = load(doc(), False)
    assert good is not None and good.n_tri == 1
    bad = [{"accessors/0/count": 2 ** 61}, {"accessors/0/count": -3}, {"accessors/0/count": 1e300}, {"accessors/0/count": 2.5},
           {"bufferViews/0/byteOffset": -8}, {"bufferViews/0/byteOffset": 2 ** 64}, {"bufferViews/0/byteOffset": 40},
           {"accessors/0/byteOffset": 2 ** 63}, {"accessors/0/byteOffset": 30}, {"bufferViews/0/byteStride": 4},
           {"bufferViews/0/byteStride": 2 ** 62}, {"bufferViews/0/byteLength": 10 ** 9}, {"accessors/1/count": 2 ** 40},
           {"bufferViews/1/byteOffset": 2 ** 53 + 2}]
    for over in bad:
        for quirks in (False, True):
            s = load(doc(**over), quirks)
            assert s is None or s.n_tri <= 1, over      # either refused, or (a field the reference's rules ignore) loaded in bounds
        assert load(doc(**over), False) is None, over


# ---- bvhNode objects nested in another bvhNode's list (bvh.h:14 allows any hittable there) -----------------------------------
def _nested_map(n_small):
    """primitive id of scene 11 / 12 -> id of config 1: [ground, big1..3 | small...] -> [ground, small..., big1..3]"""
    m = np.empty(n_small + 4, np.int64)
    m[0] = 0
    m[1:4] = n_small + 1 + np.arange(3)
    m[4:] = 1 + np.arange(n_small)
    return m


def _nested_rays(scenes, port):
    import rayset
    s1 = scenes["c1_shirley"]
    h = port.from_scene(s1)
    try:
        pin = rayset.pinhole(s1, step=2)
        want = port.trace(h, pin)
        rays = np.concatenate([pin, rayset.secondary(want, 20000, 3), rayset.random_rays(10000, 4)])
        return rays, port.trace(h, rays)
    finally:
        port.free(h)


@pytest.mark.parametrize("which", [11, 12])
def test_nested_bvh_nodes_flatten_to_the_same_hits(which, scenes, port, tmp_path):
    """Scenes 11 / 12 = config 1's spheres (same draws, same parameters) with the small spheres in a bvhNode of their own inside
    the outer bvhNode's list (12: the outer list holds one bvhNode only - a span of one, hit twice as bvh.h:67-69 does).  The
    flattening splices the nested tree in where the builder put its root box; a scene of spheres returns the nearest hit whatever
    the tree (sphere::hit honours tMax), so the hits must be config 1's, primitive for primitive."""
    s, draws = host.reference_scene(which, str(tmp_path))
    s1 = scenes["c1_shirley"]
    n_small = len(s1.sph_r) - 4
    assert len(s.sph_r) == len(s1.sph_r) and s.n_tri == 0
    m = _nested_map(n_small)
    assert np.array_equal(s.sph_c0[np.argsort(m)], s1.sph_c0) and np.array_equal(s.sph_r[np.argsort(m)], s1.sph_r)
    # a tree: every node reachable once, leaves cover every sphere (12: the nested copy is traversed twice -> twice)
    left, right = s.nodes["left"], s.nodes["right"]
    leaves = np.concatenate([~left[left < 0], ~right[(right < 0) & (right != np.iinfo(np.int32).min)]])
    cnt = np.bincount(leaves, minlength=len(s.sph_r))
    assert (cnt == (1 if which == 11 else 2)).all()
    rays, want = _nested_rays(scenes, port)
    h = port.from_scene(s)
    try:
        got = port.trace(h, rays)
    finally:
        port.free(h)
    hit = want["prim"] >= 0
    assert ((got["prim"] >= 0) == hit).all()
    assert (m[got["prim"][hit]] == want["prim"][hit]).all()
    assert (got["t"][hit].view(np.uint32) == want["t"][hit].view(np.uint32)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("which", [11, 12])
def test_nested_bvh_nodes_on_the_gpu(which, device, scenes, port, tmp_path):
    """the same scenes on the device: srt_trace_rays, the wavefront's extend kernel and a frame against the port"""
    s, _ = host.reference_scene(which, str(tmp_path))
    rays, _ = _nested_rays(scenes, port)
    h = port.from_scene(s)
    ds = device.upload(s)
    try:
        want = port.trace(h, rays)
        got = ds.trace(rays)
        assert (got["prim"] == want["prim"]).all()
        prim, t = ds.trace_wavefront(rays)
        assert (prim == want["prim"]).all()
        hit = want["prim"] >= 0
        assert (t[hit].view(np.uint32) == want["t"][hit].view(np.uint32)).all()
        s1 = scenes["c1_shirley"]
        kw = dict(width=160, height=90, spp=8, max_bounce=6, seed=3)
        P = s1.render_params(**kw)
        cam = s1.camera_struct()
        a, _, sa = ds.render(P, camera=cam, want_rgba=False)
        b, _, sb = ds.render(s1.render_params(variant=2, **kw), camera=cam, want_rgba=False)
        assert sa.paths == sb.paths == 160 * 90 * 8 and sa.segments == sb.segments
        fin = np.isfinite(a).all(axis=2)
        assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, float(np.abs(a[fin]).max()))
    finally:
        ds.close()
        port.free(h)
