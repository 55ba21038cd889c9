"""CPU tests of the CHECKER: the plain-C oracle (oracle/srt_oracle.c) against golden vectors generated from
the compiled reference (oracle/make_golden.py), bit for bit; and, where oracle/_ref exists (this container),
against the compiled reference live on larger seeded sets.  This is what pins the oracle."""
import numpy as np
import pytest

import rayset
from conftest import SCENES, psnr_u8, same_bits, tex_oob_mask
from test_gpu_parity import kat_rvec, SETS


@pytest.fixture(scope="module")
def port_scenes(port, scenes):
    out = {n: port.from_scene(s) for n, s in scenes.items()}
    yield out
    for h in out.values():
        port.free(h)


@pytest.mark.parametrize("name", SCENES)
def test_port_hits_bit_exact_vs_golden(name, port, port_scenes, golden):
    g = golden.kat(name)
    for k in SETS:
        got = port.trace(port_scenes[name], g["rays_" + k])
        assert got.tobytes() == g["hits_" + k].tobytes(), k
    got = port.trace(port_scenes[name], g["rays_bounded"], 0.05, 0.9)
    assert got.tobytes() == g["hits_bounded"].tobytes()
    assert port.count_tests(port_scenes[name], g["rays_secondary"]) == tuple(int(x) for x in g["count_secondary"])


@pytest.mark.parametrize("name", SCENES)
def test_port_shade_bit_exact_vs_golden(name, port, port_scenes, scenes, golden):
    g = golden.kat(name)
    att, emit, out, sc = port.shade_kat(port_scenes[name], g["shade_rays"], g["shade_hits"], kat_rvec(g, scenes[name]))
    assert (sc == g["shade_scattered"]).all()
    assert same_bits(emit, g["shade_emit"])
    s = sc == 1
    assert same_bits(att[s], g["shade_att"][s])
    assert out[s].tobytes() == g["shade_out"][s].tobytes()
    kinds = set(scenes[name].mat_kind[g["shade_hits"]["mat"]].tolist())
    assert 0 in kinds      # every scene exercises the PBR path


@pytest.mark.parametrize("name", SCENES)
def test_port_textures_bit_exact_vs_golden(name, port, port_scenes, scenes, golden):
    g = golden.kat(name)
    t = 0
    while "tex_uvp_%d" % t in g.files:
        uvp = g["tex_uvp_%d" % t]
        ok = ~tex_oob_mask(scenes[name], t, uvp)
        assert same_bits(port.texture_kat(port_scenes[name], t, uvp)[ok], g["tex_val_%d" % t][ok]), t
        t += 1


@pytest.mark.parametrize("name", SCENES)
def test_port_getray_bit_exact_vs_golden(name, port, scenes, golden):
    g = golden.kat(name)
    got = port.getray_kat(scenes[name].camera, g["getray_in"])
    assert got.tobytes() == g["getray_out"].tobytes()


def test_port_aabb_bit_exact_vs_golden(port, golden):
    g = golden.aabb()
    got = port.aabb_kat(g["rays"], g["boxes"], g["tminmax"])
    assert (got == g["hit"]).all()
    assert 0.05 < g["hit"].mean() < 0.95


def test_port_resolve_vs_golden(port, golden):
    g = golden.image("c2_cube")
    assert (port.resolve(g["sum_a"], int(g["spp"])) == g["rgba_a"]).all()


@pytest.mark.parametrize("name", ["c1_shirley", "c2_cube", "c3s_globe"])
def test_port_frame_vs_reference_image(name, port, port_scenes, scenes, golden):
    """The only place the port differs from the reference is the random stream (Philox + direct sampling vs
    mt19937 + rejection): its frames must sit inside the reference's own noise."""
    g = golden.image(name)
    s = scenes[name]
    H, W, _ = g["sum_a"].shape
    spp = int(g["spp"])
    P = s.render_params(width=W, height=H, spp=spp, max_bounce=int(g["depth"]), seed=9)
    got, st = port.render(port_scenes[name], s, P, nthreads=8)
    rgba = port.resolve(got, spp)
    ok = np.isfinite(g["sum_a"]).all(axis=2) & np.isfinite(g["sum_b"]).all(axis=2) & np.isfinite(got).all(axis=2)
    floor = psnr_u8(g["rgba_a"], g["rgba_b"], ok)
    mine = 0.5 * (psnr_u8(rgba, g["rgba_a"], ok) + psnr_u8(rgba, g["rgba_b"], ok))
    assert mine >= floor - 0.5, "PSNR %.2f dB vs reference noise floor %.2f dB" % (mine, floor)
    ref_spp = np.mean(g["segments"]) / (W * H * spp)
    assert abs(st.segments / st.paths - ref_spp) <= 0.01 * ref_spp
    ra, rb = g["sum_a"][ok].sum(), g["sum_b"][ok].sum()
    assert abs(got[ok].sum() - 0.5 * (ra + rb)) <= 0.02 * ra + 2 * abs(ra - rb)


def test_port_sample_split_additive(port, port_scenes, scenes):
    s = scenes["c2_cube"]
    kw = dict(width=96, height=54, spp=6, seed=3)
    full, st = port.render(port_scenes["c2_cube"], s, s.render_params(**kw), nthreads=4)
    a, sa = port.render(port_scenes["c2_cube"], s, s.render_params(sample_begin=0, sample_count=2, **kw), nthreads=2)
    b, sb = port.render(port_scenes["c2_cube"], s, s.render_params(sample_begin=2, sample_count=4, **kw), nthreads=3)
    assert sa.segments + sb.segments == st.segments
    assert np.abs((a + b) - full).max() <= 1e-5 * np.abs(full).max()


def test_port_rotated_grid_vs_golden(port, golden, port_scenes, scenes):
    """sampler 1: the port's in-pixel positions equal, bit for bit, what the reference's own statements (main.cpp:181,205-207,
    run through Eigen's AngleAxisf by oracle/ref_oracle.cpp) gave for samples 0..63; the frame it renders is deterministic,
    differs from the jittered one and splits additively over sample ranges like sampler 0."""
    pos = golden("sampler")["pos"]
    for s in range(len(pos)):
        assert same_bits(port.rotated_grid(s), pos[s]), s
    assert np.abs(np.hypot(pos[:, 0], pos[:, 1]) - 0.8).max() < 1e-6
    sc = scenes["c2_cube"]
    h = port_scenes["c2_cube"]
    P = sc.render_params(width=64, height=36, spp=8, seed=3, sampler=1)
    a, _ = port.render(h, sc, P, nthreads=4)
    b, _ = port.render(h, sc, P, nthreads=1)
    assert same_bits(a, b)
    j, _ = port.render(h, sc, sc.render_params(width=64, height=36, spp=8, seed=3), nthreads=4)
    assert np.abs(a - j).max() > 0.05
    lo, _ = port.render(h, sc, sc.render_params(width=64, height=36, spp=8, seed=3, sampler=1, sample_begin=0, sample_count=5), nthreads=4)
    hi, _ = port.render(h, sc, sc.render_params(width=64, height=36, spp=8, seed=3, sampler=1, sample_begin=5, sample_count=3), nthreads=4)
    assert np.allclose(lo + hi, a, rtol=1e-5, atol=1e-6)


def test_ref_rotated_grid_live(ref, golden):
    pos = golden("sampler")["pos"]
    for s in range(len(pos)):
        assert same_bits(ref.rotated_grid(s), pos[s])


def test_port_smooth_normals_and_packed_mr(port, scenes, ext_scene):
    """The two non-reference options of srt.h (tri_n0..2, SRT_MATF_PACKED_MR) in the port, against numpy restatements:
    the shading normal is unit(sum r_k n_k) with the inverse-distance weights of model.h:158-166, ids / t / uv / tangent basis
    are untouched; a packed map gives metalness = B/255 * factor, roughness = G/255 * factor - checked against a scene whose
    material carries those two numbers as scalars, texel by texel."""
    base = scenes["c3_masterchief"]
    rays = rayset.pinhole(base, step=3)
    hb, he = port.from_scene(base), port.from_scene(ext_scene)
    a, b = port.trace(hb, rays), port.trace(he, rays)
    for f in ("prim", "mat", "front_face", "t", "p", "uv", "tangent", "bitangent"):
        assert same_bits(a[f], b[f]), f
    tri = (a["prim"] >= 0) & (a["prim"] < base.n_tri)
    assert tri.sum() > 500
    assert same_bits(a["normal"][~tri], b["normal"][~tri])
    F = np.float32
    i = a["prim"][tri]
    p = a["p"][tri]
    w = [F(1) / np.sqrt((((p - v[i]) ** 2).astype(F)).sum(axis=1, dtype=np.float64)).astype(F) for v in (base.tri_v0, base.tri_v1, base.tri_v2)]
    den = (w[0] + w[1] + w[2]).astype(F)
    r = [(x / den).astype(F) for x in w]
    n = r[0][:, None] * ext_scene.tri_n0[i] + r[1][:, None] * ext_scene.tri_n1[i] + r[2][:, None] * ext_scene.tri_n2[i]
    n = n / np.linalg.norm(n, axis=1, keepdims=True)
    assert np.abs(b["normal"][tri] - n).max() < 2e-5
    assert np.abs(b["normal"][tri] - a["normal"][tri]).max() > 0.05
    # packed metallic-roughness: shade the flat-normal hits with the packed map, then with per-hit scalars
    import copy
    flat = copy.copy(ext_scene)
    flat.tri_n0 = flat.tri_n1 = flat.tri_n2 = None
    hf = port.from_scene(flat)
    sel = np.flatnonzero(tri)[:400]
    rng = np.random.default_rng(11)
    rv = rng.normal(size=(len(sel), 3))
    xi = np.concatenate([rv / np.linalg.norm(rv, axis=1, keepdims=True), rng.random((len(sel), 1))], axis=1).astype(F)
    att, _, out, sc = port.shade_kat(hf, rays[sel], a[sel], xi)
    att0, _, _, _ = port.shade_kat(hb, rays[sel], a[sel], xi)
    assert np.abs(att - att0).max() > 1e-3
    for k in range(0, len(sel), 40):
        hit = a[sel[k]]
        m = int(hit["mat"])
        assert flat.mat_flags[m] == 1
        tex = int(flat.mat_tex[m, 2])
        texel = port.texture_kat(hf, tex, np.array([[hit["uv"][0], hit["uv"][1], 0, 0, 0]], F))[0]
        one = copy.copy(flat)
        one.mat_tex = flat.mat_tex.copy(); one.mat_params = flat.mat_params.copy(); one.mat_flags = None
        one.mat_tex[m, 2] = -1
        one.mat_params[m, 0] = min(max(F(texel[2]) / F(255) * flat.mat_params[m, 0], F(0)), F(1))
        one.mat_params[m, 1] = min(max(F(texel[1]) / F(255) * flat.mat_params[m, 1], F(0)), F(1))
        h1 = port.from_scene(one)
        att1, _, out1, _ = port.shade_kat(h1, rays[sel[k]:sel[k] + 1], a[sel[k]:sel[k] + 1], xi[k:k + 1])
        port.free(h1)
        assert same_bits(att1[0], att[k]) and same_bits(out1, out[k:k + 1])
    for h in (hb, he, hf):
        port.free(h)


def test_philox_known_answers(port):
    """Philox4x32-10 known-answer vectors of the Random123 distribution (kat_vectors): counter, key -> output.
    so_philox maps (seed, pixel, sample, block) to key=(seed lo, hi), counter=(pixel, sample, block, 0)."""
    got = port.philox(0, 0, 0, 0)
    assert [hex(int(x)) for x in got] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]


# ---- live against the compiled reference (this container only) -----------------------------------------
@pytest.mark.parametrize("name", SCENES)
def test_port_vs_reference_live(name, port, port_scenes, ref, scenes):
    s = scenes[name]
    h = ref.from_scene(s)
    pin = rayset.pinhole(s, step=2)
    want = ref.trace(h, pin)
    assert port.trace(port_scenes[name], pin).tobytes() == want.tobytes()
    for rays in (rayset.secondary(want, 30000, 21), rayset.random_rays(30000, 22), rayset.adversarial(s, 20000, 23)):
        assert port.trace(port_scenes[name], rays).tobytes() == ref.trace(h, rays).tobytes()
    ref.free(h)


def test_reference_from_desc_equals_native(ref, scenes):
    """The scene files are faithful: the reference rebuilt from the flattened description answers exactly
    like the reference built by its own loader (gltfLoad + bvhNode ctor)."""
    s = scenes["c3_masterchief"]
    a, b = ref.native(3), ref.from_scene(s)
    rays = np.concatenate([rayset.pinhole(s, step=3), rayset.random_rays(20000, 31)])
    assert ref.trace(a, rays).tobytes() == ref.trace(b, rays).tobytes()
    ref.free(a)
    ref.free(b)


def test_port_aabb_adversarial_vs_reference(port, ref):
    """aabb::hit on the generated threshold set (rayset.aabb_adversarial) that the GPU predicates are held to: the C port's
    decisions equal the compiled reference's, so the GPU box can use the port as the checker for this set."""
    import rayset
    for seed in (11, 12, 13):
        rays, boxes, tmm = rayset.aabb_adversarial(300000, seed)
        assert (port.aabb_kat(rays, boxes, tmm) == ref.aabb_kat(rays, boxes, tmm)).all()


def test_port_vertex_tangents(port, ext_scene, tan_scene):
    """srt_scene_desc::tri_tan0..2 (glTF TANGENT, never read by the reference) in the port, against a numpy restatement:
    tangent = unit(sum r_k t_k) with the inverse-distance weights of model.h:158-166, bitangent = w0 * unit(normal x tangent);
    triangles whose vertices carry w = 0 keep calcTangentBasis' per-face basis bit for bit; nothing else of the record moves."""
    rays = rayset.pinhole(ext_scene, step=3)
    he, ht = port.from_scene(ext_scene), port.from_scene(tan_scene)
    try:
        a, b = port.trace(he, rays), port.trace(ht, rays)
    finally:
        port.free(he); port.free(ht)
    for f in ("prim", "mat", "front_face", "t", "p", "uv", "normal"):
        assert same_bits(a[f], b[f]), f
    tri = (a["prim"] >= 0) & (a["prim"] < ext_scene.n_tri)
    i = a["prim"][tri]
    w0 = tan_scene.tri_tan0[i][:, 3]
    keep = w0 == 0
    assert keep.sum() > 100 and (~keep).sum() > 300
    assert same_bits(a["tangent"][~tri], b["tangent"][~tri]) and same_bits(a["bitangent"][~tri], b["bitangent"][~tri])
    assert same_bits(a["tangent"][tri][keep], b["tangent"][tri][keep]) and same_bits(a["bitangent"][tri][keep], b["bitangent"][tri][keep])
    F = np.float32
    p = a["p"][tri]
    wt = [F(1) / np.sqrt((((p - v[i]) ** 2).astype(F)).sum(axis=1, dtype=np.float64)).astype(F) for v in (ext_scene.tri_v0, ext_scene.tri_v1, ext_scene.tri_v2)]
    den = (wt[0] + wt[1] + wt[2]).astype(F)
    r = [(x / den).astype(F) for x in wt]
    T = r[0][:, None] * tan_scene.tri_tan0[i][:, :3] + r[1][:, None] * tan_scene.tri_tan1[i][:, :3] + r[2][:, None] * tan_scene.tri_tan2[i][:, :3]
    with np.errstate(invalid="ignore", divide="ignore"):        # (rows without tangents are masked out below)
        T = T / np.linalg.norm(T, axis=1, keepdims=True)
        B = np.cross(b["normal"][tri].astype(np.float64), T)
        B = w0[:, None] * B / np.linalg.norm(B, axis=1, keepdims=True)
    use = ~keep & np.isfinite(T).all(axis=1) & np.isfinite(B).all(axis=1)
    assert np.abs(b["tangent"][tri][use] - T[use]).max() < 2e-5
    assert np.abs(b["bitangent"][tri][use] - B[use]).max() < 5e-5
    assert np.abs(b["tangent"][tri][use] - a["tangent"][tri][use]).max() > 0.05        # and it is not the per-face basis
