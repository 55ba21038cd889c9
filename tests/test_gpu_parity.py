"""GPU parity tests: the CUDA path, called through the C ABI (libsrt_b200.so), against
(1) golden vectors generated from the compiled reference (tests/golden, oracle/make_golden.py) and
(2) the plain-C oracle run live on the same seeded inputs.

Bars (north_star): hit primitive id, miss/hit and front-face BIT-EXACT; t / p / normal / uv within
1e-5 relative (in practice they are bit-identical except where acosf/atan2f differ from glibc by ulps);
shading functions within 1e-4 relative; frames within the reference's own noise floor - 0.5 dB.
"""
import numpy as np
import pytest

import rayset
from conftest import SCENES, psnr_u8, rel_err, tex_oob_mask

pytestmark = pytest.mark.gpu

SETS = ["pinhole", "secondary", "random", "adversarial"]


def check_hits(got, want, tol=1e-5):
    assert (got["prim"] == want["prim"]).all(), "primitive ids differ at %s" % np.flatnonzero(got["prim"] != want["prim"])[:10]
    assert (got["mat"] == want["mat"]).all()
    assert (got["front_face"] == want["front_face"]).all()
    h = want["prim"] >= 0
    for f in ("t", "p", "normal", "tangent", "bitangent", "uv"):
        a, b = got[f][h], want[f][h]
        nan = np.isnan(b)
        assert (np.isnan(a) == nan).all(), f
        # relative 1e-5; components of unit vectors and sphere uv (= (atan2f+pi)/2pi, acosf/pi: cancellation
        # against pi leaves an absolute accuracy of ulp(pi)) get an absolute floor of 1e-7
        floor = 1e-7 if f in ("uv", "tangent", "bitangent", "normal") else 1e-12
        with np.errstate(invalid="ignore"):
            err = np.where(nan, 0, (np.abs(a.astype(np.float64) - b) - floor) / np.maximum(np.abs(b), 1e-30))
        assert err.max() <= tol, "%s max rel err %g" % (f, err.max())


@pytest.mark.parametrize("name", SCENES)
def test_trace_vs_reference_golden(name, dev_scenes, golden):
    g = golden.kat(name)
    for k in SETS:
        check_hits(dev_scenes[name].trace(g["rays_" + k]), g["hits_" + k])
    check_hits(dev_scenes[name].trace(g["rays_bounded"], 0.05, 0.9), g["hits_bounded"])


@pytest.mark.parametrize("name", SCENES)
def test_trace_vs_port_live(name, dev_scenes, scenes, port):
    s = scenes[name]
    h = port.from_scene(s)
    pin = rayset.pinhole(s)
    want = port.trace(h, pin)
    check_hits(dev_scenes[name].trace(pin), want)
    for rays in (rayset.secondary(want, 60000, 5), rayset.random_rays(60000, 6), rayset.adversarial(s, 40000, 7)):
        check_hits(dev_scenes[name].trace(rays), port.trace(h, rays))
    port.free(h)


@pytest.mark.parametrize("name", SCENES)
def test_trace_counters_match_reference(name, device, dev_scenes, golden):
    """Same traversal order => same number of box tests and leaf tests as the reference's recursion
    (span-1 nodes visit their leaf once instead of twice: leaf count may only be smaller)."""
    g = golden.kat(name)
    device.set_counters(True)
    try:
        dev_scenes[name].trace(g["rays_secondary"])
        c = device.last_counters()
    finally:
        device.set_counters(False)
    box, leaf = (int(x) for x in g["count_secondary"])
    assert c["node_visits"] == box
    assert c["leaf_tests"] <= leaf
    assert c["exact_box_tests"] < 0.02 * box     # the filtered predicate decides almost everything


def unit_rows(p):
    F = np.float32
    ls = ((p[:, 0] * p[:, 0] + p[:, 1] * p[:, 1]).astype(F) + p[:, 2] * p[:, 2]).astype(F)
    ln = np.sqrt(ls).astype(F)
    return (p / ln[:, None]).astype(F)


def kat_rvec(g, scene):
    """The random vector the reference drew from the injected numbers: p = -1 + 2*xi (randomFloat(-1,1)),
    unit vector for PBR (randomUnitVector), p itself for METAL (randomInUnitSphere)."""
    F = np.float32
    xi = g["shade_xi"]
    p = (F(-1) + F(2) * xi[:, :3]).astype(F)
    kind = scene.mat_kind[g["shade_hits"]["mat"]]
    rv = np.where((kind == 0)[:, None], unit_rows(p), p).astype(F)
    return np.concatenate([rv, xi[:, 3:4]], axis=1).astype(F)


@pytest.mark.parametrize("name", SCENES)
def test_shade_vs_reference_golden(name, dev_scenes, scenes, golden):
    g = golden.kat(name)
    att, emit, out, sc = dev_scenes[name].shade_kat(g["shade_rays"], g["shade_hits"], kat_rvec(g, scenes[name]))
    assert (sc == g["shade_scattered"]).mean() > 0.9995     # metal's dir.n > 0 can flip on an ulp
    ok = sc == g["shade_scattered"]
    assert rel_err(emit, g["shade_emit"]).max() <= 1e-6
    s = ok & (sc == 1)
    scale = np.maximum(np.abs(g["shade_att"][s]), 1e-4)
    assert (np.abs(att[s] - g["shade_att"][s]) / scale).max() <= 1e-4
    assert np.abs(out["dir"][s] - g["shade_out"]["dir"][s]).max() <= 2e-6 * max(1.0, np.abs(g["shade_out"]["dir"][s]).max())
    assert (out["o"][s] == g["shade_out"]["o"][s]).all()
    assert (out["time"][s] == g["shade_out"]["time"][s]).all()


@pytest.mark.parametrize("name", SCENES)
def test_textures_vs_reference_golden(name, dev_scenes, scenes, golden):
    g = golden.kat(name)
    t = 0
    while "tex_uvp_%d" % t in g.files:
        ok = ~tex_oob_mask(scenes[name], t, g["tex_uvp_%d" % t])
        got = dev_scenes[name].texture_kat(t, g["tex_uvp_%d" % t])[ok]
        want = g["tex_val_%d" % t][ok]
        bad = np.abs(got - want).max(axis=1) > 1e-4 * max(1.0, np.abs(want).max())
        assert not bad.any(), "texture %d: %d lookups differ, first %s got %s want %s" % (
            t, bad.sum(), g["tex_uvp_%d" % t][ok][bad][0], got[bad][0], want[bad][0])
        t += 1
    assert t > 0


@pytest.mark.parametrize("name", ["c1_shirley", "c3_masterchief"])
def test_getray_vs_reference_golden(name, device, scenes, golden):
    g = golden.kat(name)
    got = device.getray_kat(scenes[name].camera_struct(), g["getray_in"])
    want = g["getray_out"]
    for f in ("o", "dir", "time"):
        assert (got[f].view(np.uint32) == want[f].view(np.uint32)).all(), f


def test_primary_rays_vs_port(device, scenes, port):
    """Philox stream + direct disk sampling + pixel jitter: GPU and C oracle agree (sincosf vs glibc: ulps)."""
    s = scenes["c3_masterchief"]
    P = s.render_params(seed=77)
    rng = np.random.default_rng(3)
    n = 5000
    xy = np.stack([rng.integers(0, P.width, n), rng.integers(0, P.height, n)], 1).astype(np.int32)
    smp = rng.integers(0, 5000, n).astype(np.int32)
    got = device.primary_rays(s.camera_struct(), P, xy, smp)
    F = np.float32
    inp = np.zeros((n, 5), F)
    for i in range(n):
        a = port.philox(77, int(xy[i, 1]) * P.width + int(xy[i, 0]), int(smp[i]), 0)
        c = port.philox(77, int(xy[i, 1]) * P.width + int(xy[i, 0]), int(smp[i]), 1)
        u = (a >> 8).astype(F) * F(1.0 / 16777216.0)
        inp[i, 0] = (F(xy[i, 0]) + u[0]) / F(P.width - 1)
        inp[i, 1] = (F(P.height - xy[i, 1]) + u[1]) / F(P.height - 1)
        r, phi = np.sqrt(u[2]), F(2.0) * F(np.pi) * u[3]
        inp[i, 2], inp[i, 3] = r * np.cos(phi, dtype=F), r * np.sin(phi, dtype=F)
        inp[i, 4] = (c[0] >> 8).astype(F) * F(1.0 / 16777216.0)
    want = port.getray_kat(s.camera, inp)
    assert np.abs(got["o"] - want["o"]).max() <= 1e-6
    assert np.abs(got["dir"] - want["dir"]).max() <= 1e-5
    assert (got["time"] == want["time"]).all()


def test_primary_rays_rotated_grid(device, scenes, port, golden):
    """sampler 1 (main.cpp:205-207,214-215): the in-pixel position of sample s is the golden vector generated by running
    the reference's own Eigen statements; the rays built from it agree with the port's camera."""
    s = scenes["c3_masterchief"]
    P = s.render_params(seed=77, sampler=1)
    pos = golden("sampler")["pos"]
    rng = np.random.default_rng(4)
    n = 3000
    xy = np.stack([rng.integers(0, P.width, n), rng.integers(0, P.height, n)], 1).astype(np.int32)
    smp = rng.integers(0, len(pos), n).astype(np.int32)
    got = device.primary_rays(s.camera_struct(), P, xy, smp)
    F = np.float32
    inp = np.zeros((n, 5), F)
    for i in range(n):
        a = port.philox(77, int(xy[i, 1]) * P.width + int(xy[i, 0]), int(smp[i]), 0)
        c = port.philox(77, int(xy[i, 1]) * P.width + int(xy[i, 0]), int(smp[i]), 1)
        u = (a >> 8).astype(F) * F(1.0 / 16777216.0)
        inp[i, 0] = (F(xy[i, 0]) + pos[smp[i], 0]) / F(P.width - 1)
        inp[i, 1] = (F(P.height - xy[i, 1]) + pos[smp[i], 1]) / F(P.height - 1)
        r, phi = np.sqrt(u[2]), F(2.0) * F(np.pi) * u[3]
        inp[i, 2], inp[i, 3] = r * np.cos(phi, dtype=F), r * np.sin(phi, dtype=F)
        inp[i, 4] = (c[0] >> 8).astype(F) * F(1.0 / 16777216.0)
    want = port.getray_kat(s.camera, inp)
    assert np.abs(got["o"] - want["o"]).max() <= 1e-6
    assert np.abs(got["dir"] - want["dir"]).max() <= 1e-5
    assert (got["time"] == want["time"]).all()
    # with a pin-hole camera the ray depends on the sampler alone: bit-exact
    cam = s.camera_struct()
    cam24 = np.array(s.camera, F).copy()
    pin = scenes["c2_cube"]                      # aperture 0
    Pp = pin.render_params(seed=1, sampler=1)
    xy = np.stack([rng.integers(0, Pp.width, n), rng.integers(0, Pp.height, n)], 1).astype(np.int32)
    got = device.primary_rays(pin.camera_struct(), Pp, xy, smp)
    inp = np.zeros((n, 5), F)
    inp[:, 0] = (xy[:, 0].astype(F) + pos[smp, 0]) / F(Pp.width - 1)
    inp[:, 1] = ((Pp.height - xy[:, 1]).astype(F) + pos[smp, 1]) / F(Pp.height - 1)
    want = port.getray_kat(pin.camera, inp)
    assert (got["dir"].view(np.uint32) == want["dir"].view(np.uint32)).all()


# ---- frames ------------------------------------------------------------------------------------------
SMALL = {"c1_shirley": (200, 112, 8), "c2_cube": (160, 90, 8), "c3_masterchief": (213, 120, 8), "c3s_globe": (213, 120, 8)}


@pytest.mark.parametrize("name", SCENES)
def test_frame_vs_port_same_stream(name, dev_scenes, scenes, port):
    """Same Philox stream on both sides: per-pixel sums agree except where an ulp of sinf/acosf/... flips
    a discrete decision somewhere along a path; segment counts agree to 1e-3."""
    s = scenes[name]
    W, H, spp = SMALL[name]
    P = s.render_params(width=W, height=H, spp=spp, seed=5)
    h = port.from_scene(s)
    want, st_w = port.render(h, s, P, nthreads=8)
    port.free(h)
    got, rgba, st = dev_scenes[name].render(P)
    assert st.paths == W * H * spp == st_w.paths
    assert abs(int(st.segments) - int(st_w.segments)) <= 1e-3 * st_w.segments
    fin = np.isfinite(want).all(axis=2) & np.isfinite(got).all(axis=2)
    assert fin.mean() > 0.999
    err = np.abs(got - want).max(axis=2) / np.maximum(np.abs(want).max(axis=2), 1e-3)
    assert (err[fin] > 1e-3).mean() < 0.01, "fraction of differing pixels %g" % (err[fin] > 1e-3).mean()
    assert abs(got[fin].sum() - want[fin].sum()) <= 2e-3 * want[fin].sum()
    assert (rgba[..., 3] == 255).all()


@pytest.mark.parametrize("variant", [1, 2, 3])
def test_frame_rotated_grid_vs_port(variant, dev_scenes, scenes, port):
    """sampler 1 in every render variant against the port on the same stream, 4 samples (one turn of the grid) and 9."""
    s = scenes["c3_masterchief"]
    h = port.from_scene(s)
    for spp, begin in ((4, 0), (9, 3)):
        P = s.render_params(width=213, height=120, spp=begin + spp, sample_begin=begin, sample_count=spp, seed=5, sampler=1,
                            variant=variant)
        want, st_w = port.render(h, s, P, nthreads=8)
        got, _, st = dev_scenes["c3_masterchief"].render(P, want_rgba=False)
        assert st.variant == variant and st.paths == st_w.paths
        fin = np.isfinite(want).all(axis=2) & np.isfinite(got).all(axis=2)
        err = np.abs(got - want).max(axis=2) / np.maximum(np.abs(want).max(axis=2), 1e-3)
        assert (err[fin] > 1e-3).mean() < 0.01
        assert abs(got[fin].sum() - want[fin].sum()) <= 2e-3 * want[fin].sum()
        P0 = s.render_params(width=213, height=120, spp=begin + spp, sample_begin=begin, sample_count=spp, seed=5, variant=variant)
        other, _, _ = dev_scenes["c3_masterchief"].render(P0, want_rgba=False)
        assert np.abs(other - got).max() > 0.1            # the sampler does change the frame
    port.free(h)
    Pbad = s.render_params(sampler=2)
    with pytest.raises(Exception):
        dev_scenes["c3_masterchief"].render(Pbad)


def test_smooth_normals_and_packed_mr_vs_port(device, ext_scene, port):
    """srt.h's two non-reference options (tri_n0..2, SRT_MATF_PACKED_MR): hit records, shading and whole frames of the CUDA
    path against the C port on the same inputs."""
    import rayset
    ds = device.upload(ext_scene)
    h = port.from_scene(ext_scene)
    try:
        rays = rayset.pinhole(ext_scene, step=2)
        want = port.trace(h, rays)
        rays = np.concatenate([rays, rayset.secondary(want, 20000, 5)])
        want = port.trace(h, rays)
        got = ds.trace(rays)
        assert (got["prim"] == want["prim"]).all() and (got["front_face"] == want["front_face"]).all()
        hit = want["prim"] >= 0
        assert np.abs(got["normal"][hit] - want["normal"][hit]).max() <= 1e-5
        assert rel_err(got["t"][hit], want["t"][hit]).max() <= 1e-5
        tri = hit & (want["prim"] < ext_scene.n_tri)
        sel = np.flatnonzero(tri)[::7][:5000]
        rng = np.random.default_rng(12)
        rv = rng.normal(size=(len(sel), 3))
        xi = np.concatenate([rv / np.linalg.norm(rv, axis=1, keepdims=True), rng.random((len(sel), 1))], axis=1).astype(np.float32)
        att_w, _, out_w, sc_w = port.shade_kat(h, rays[sel], want[sel], xi)
        att_g, _, out_g, sc_g = ds.shade_kat(rays[sel], want[sel], xi)
        assert (sc_w == sc_g).all()
        assert (np.abs(att_g - att_w) / np.maximum(np.abs(att_w), 1e-4)).max() <= 1e-4
        assert np.abs(out_g["dir"] - out_w["dir"]).max() <= 2e-6
        for variant in (1, 2, 3):
            P = ext_scene.render_params(width=213, height=120, spp=8, seed=5, variant=variant)
            want_f, st_w = port.render(h, ext_scene, P, nthreads=8)
            got_f, _, st = ds.render(P, want_rgba=False)
            assert st.variant == variant and st.paths == st_w.paths
            assert abs(int(st.segments) - int(st_w.segments)) <= 1e-3 * st_w.segments
            fin = np.isfinite(want_f).all(axis=2) & np.isfinite(got_f).all(axis=2)
            err = np.abs(got_f - want_f).max(axis=2) / np.maximum(np.abs(want_f).max(axis=2), 1e-3)
            assert (err[fin] > 1e-3).mean() < 0.01
            assert abs(got_f[fin].sum() - want_f[fin].sum()) <= 2e-3 * want_f[fin].sum()
    finally:
        port.free(h)
        ds.close()
    # all three normal arrays or none
    import ctypes
    from sexy_raytracer_b200 import device as devmod
    d = ext_scene.desc()
    d.tri_n1 = None
    hs = ctypes.c_void_p()
    assert devmod.lib().srt_scene_create(device.h, ctypes.byref(d), ctypes.byref(hs)) != 0 and not hs


def test_vertex_tangents_vs_port(device, tan_scene, port):
    """srt_scene_desc::tri_tan0..2 (glTF TANGENT; not the reference): hit records, shading of normal-mapped materials and whole
    frames of the CUDA path against the C port on the same inputs."""
    import rayset
    ds = device.upload(tan_scene)
    h = port.from_scene(tan_scene)
    try:
        rays = rayset.pinhole(tan_scene, step=2)
        want = port.trace(h, rays)
        rays = np.concatenate([rays, rayset.secondary(want, 20000, 5)])
        want = port.trace(h, rays)
        got = ds.trace(rays)
        assert (got["prim"] == want["prim"]).all()
        tri = (want["prim"] >= 0) & (want["prim"] < tan_scene.n_tri)
        fin = tri & np.isfinite(want["tangent"]).all(axis=1) & np.isfinite(want["bitangent"]).all(axis=1)
        assert fin.sum() > 1000
        assert np.abs(got["tangent"][fin] - want["tangent"][fin]).max() <= 2e-5
        assert np.abs(got["bitangent"][fin] - want["bitangent"][fin]).max() <= 2e-5
        sel = np.flatnonzero(fin)[::7][:5000]
        rng = np.random.default_rng(12)
        rv = rng.normal(size=(len(sel), 3))
        xi = np.concatenate([rv / np.linalg.norm(rv, axis=1, keepdims=True), rng.random((len(sel), 1))], axis=1).astype(np.float32)
        att_w, _, out_w, sc_w = port.shade_kat(h, rays[sel], want[sel], xi)
        att_g, _, out_g, sc_g = ds.shade_kat(rays[sel], want[sel], xi)
        assert (sc_w == sc_g).all()
        assert (np.abs(att_g - att_w) / np.maximum(np.abs(att_w), 1e-4)).max() <= 1e-4
        for variant in (1, 2):
            P = tan_scene.render_params(width=213, height=120, spp=8, seed=5, variant=variant)
            want_f, st_w = port.render(h, tan_scene, P, nthreads=8)
            got_f, _, st = ds.render(P, want_rgba=False)
            assert st.variant == variant and st.paths == st_w.paths
            assert abs(int(st.segments) - int(st_w.segments)) <= 1e-3 * st_w.segments
            fin2 = np.isfinite(want_f).all(axis=2) & np.isfinite(got_f).all(axis=2)
            err = np.abs(got_f - want_f).max(axis=2) / np.maximum(np.abs(want_f).max(axis=2), 1e-3)
            assert (err[fin2] > 1e-3).mean() < 0.01
            assert abs(got_f[fin2].sum() - want_f[fin2].sum()) <= 2e-3 * want_f[fin2].sum()
    finally:
        port.free(h)
        ds.close()
    import ctypes
    from sexy_raytracer_b200 import device as devmod
    d = tan_scene.desc()
    d.tri_tan2 = None                             # all three tangent arrays or none
    hs = ctypes.c_void_p()
    assert devmod.lib().srt_scene_create(device.h, ctypes.byref(d), ctypes.byref(hs)) != 0 and not hs


@pytest.mark.parametrize("name", SCENES)
def test_frame_vs_reference_image(name, dev_scenes, scenes, golden):
    """Converged-image criterion: PSNR(GPU, ref_A) >= PSNR(ref_A, ref_B) - 0.5 dB on RGBA8, and
    16x16-tile means of the linear sums within 2 % + noise, at the spp the reference frames were rendered with."""
    g = golden.image(name)
    s = scenes[name]
    H, W, _ = g["sum_a"].shape
    spp = int(g["spp"])
    P = s.render_params(width=W, height=H, spp=spp, max_bounce=int(g["depth"]), seed=9)
    got, rgba, st = dev_scenes[name].render(P)
    a, b = g["rgba_a"], g["rgba_b"]
    ok = np.isfinite(g["sum_a"]).all(axis=2) & np.isfinite(g["sum_b"]).all(axis=2) & np.isfinite(got).all(axis=2)
    floor = psnr_u8(a, b, ok)
    mine = 0.5 * (psnr_u8(rgba, a, ok) + psnr_u8(rgba, b, ok))
    assert mine >= floor - 0.5, "PSNR %.2f dB vs reference noise floor %.2f dB" % (mine, floor)
    # segments per path agree with the reference's own counter
    ref_spp = np.mean(g["segments"]) / (W * H * spp)
    assert abs(st.segments / st.paths - ref_spp) <= 0.01 * ref_spp
    # tile means of the linear radiance: GPU vs mean(A,B), judged against |A-B|
    def tiles(x):
        x = np.where(ok[..., None], x, 0)[: H // 16 * 16, : W // 16 * 16]
        return x.reshape(H // 16, 16, W // 16, 16, 3).mean(axis=(1, 3))
    ta, tb, tg = tiles(g["sum_a"]), tiles(g["sum_b"]), tiles(got)
    ref = 0.5 * (ta + tb)
    # unbiased => RMS(tg - ref) / RMS(ta - tb) = sqrt(1.5 / 2) = 0.87 (relative deviations, fireflies included)
    n = np.abs(ref) + 1e-3 * spp
    ratio = np.sqrt((((tg - ref) / n) ** 2).mean()) / np.sqrt((((ta - tb) / n) ** 2).mean())
    assert ratio <= 1.1, "tile-mean deviation ratio %.3f (0.87 expected for an unbiased frame)" % ratio
    assert abs(tg.sum() - ref.sum()) <= 0.02 * ref.sum() + 2 * abs(ta.sum() - tb.sum())


def test_resolve_matches_reference(device, dev_scenes, scenes, golden):
    """writeColorTarget (color.h:25-41): feeding the reference's own sums through srt_render's resolve is not
    possible from outside, so check the rule on the GPU frame: rgba == uint8(256*clamp(sqrt(sum/spp),0,.999))."""
    s = scenes["c2_cube"]
    P = s.render_params(width=160, height=90, spp=4, seed=2)
    got, rgba, _ = dev_scenes["c2_cube"].render(P)
    F = np.float32
    c = np.sqrt((got * (F(1.0) / F(4))).astype(F)).astype(F)
    want = (F(256) * np.clip(c, F(0), F(0.999))).astype(F).astype(np.int32).astype(np.uint8)
    fin = np.isfinite(got).all(axis=2)
    assert (rgba[..., :3][fin] == want[fin]).all()


def test_sample_split_is_additive(dev_scenes, scenes):
    """spp split across GPUs: samples [0,a) + [a,n) == [0,n) up to fp32 summation order."""
    s = scenes["c3_masterchief"]
    ds = dev_scenes["c3_masterchief"]
    kw = dict(width=213, height=120, spp=16, seed=4)
    full, _, st = ds.render(s.render_params(**kw))
    a, _, sa = ds.render(s.render_params(sample_begin=0, sample_count=5, **kw), want_rgba=False)
    b, _, sb = ds.render(s.render_params(sample_begin=5, sample_count=11, **kw), want_rgba=False)
    assert sa.segments + sb.segments == st.segments
    fin = np.isfinite(full).all(axis=2)
    assert np.abs((a + b)[fin] - full[fin]).max() <= 1e-4 * max(1.0, np.abs(full[fin]).max())
    again, _, _ = ds.render(s.render_params(**kw))
    assert np.abs(again[fin] - full[fin]).max() <= 1e-4 * max(1.0, np.abs(full[fin]).max())


@pytest.mark.parametrize("name", SCENES)
def test_variants_agree(name, device, dev_scenes, scenes):
    """k_render_phased (variant 3, warp-voted phases) only re-schedules lanes: same Philox stream, same per-lane
    sample order => the fp32 sums are BIT-identical to k_render_mega (variant 1), counters included."""
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=11)
    a, ra, sa = dev_scenes[name].render(s.render_params(variant=1, **kw))
    b, rb, sb = dev_scenes[name].render(s.render_params(variant=3, **kw))
    assert sa.paths == sb.paths and sa.segments == sb.segments
    assert (a.view(np.uint32) == b.view(np.uint32)).all()
    assert (ra == rb).all()
    device.set_counters(True)
    try:
        _, _, ca = dev_scenes[name].render(s.render_params(variant=1, **kw), want_rgba=False)
        k1 = device.last_counters()
        _, _, cb = dev_scenes[name].render(s.render_params(variant=3, **kw), want_rgba=False)
        k3 = device.last_counters()
    finally:
        device.set_counters(False)
    assert ca.node_visits == cb.node_visits and ca.leaf_tests == cb.leaf_tests
    assert k1["sphere_leaf_tests"] == k3["sphere_leaf_tests"] and k1["triangle_hits"] == k3["triangle_hits"]


@pytest.mark.parametrize("name", SCENES)
def test_wavefront_agrees(name, device, dev_scenes, scenes):
    """Variant 2 (extend / shade waves over a slot pool) traces the same paths as the megakernel - same Philox
    keys, same traversal - and differs only in the order a pixel's samples are summed (red.global.add)."""
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=11)
    a, ra, sa = dev_scenes[name].render(s.render_params(variant=1, **kw))
    b, rb, sb = dev_scenes[name].render(s.render_params(variant=2, **kw))
    assert sa.paths == sb.paths and sa.segments == sb.segments
    assert sb.launches >= 3
    fin = np.isfinite(a).all(axis=2)
    assert (np.isfinite(b).all(axis=2) == fin).all()
    assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert (np.abs(ra.astype(int) - rb.astype(int)) <= 1).all()
    device.set_counters(True)
    try:
        _, _, ca = dev_scenes[name].render(s.render_params(variant=1, **kw), want_rgba=False)
        k1 = device.last_counters()
        _, _, cb = dev_scenes[name].render(s.render_params(variant=2, **kw), want_rgba=False)
        k2 = device.last_counters()
    finally:
        device.set_counters(False)
    # k_wf_extend2 culls inner nodes conservatively over a tree with fewer inner nodes: other boxes, never other primitives
    assert cb.node_visits <= 1.02 * ca.node_visits + 64 and ca.leaf_tests == cb.leaf_tests
    assert k1["sphere_leaf_tests"] == k2["sphere_leaf_tests"] and k1["triangle_hits"] == k2["triangle_hits"]
    # a slot pool smaller than the frame (many waves) and a sample sub-range give the same frame
    import os
    os.environ["SRT_WF"] = "4096,256,12"
    try:
        c, _, sc_ = dev_scenes[name].render(s.render_params(variant=2, **kw), want_rgba=False)
    finally:
        del os.environ["SRT_WF"]
    assert sc_.segments == sa.segments and sc_.launches > sb.launches
    assert np.abs(a[fin] - c[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())


def test_full_size_properties(dev_scenes, scenes):
    """BASELINE config 4 geometry (1280x720) at reduced spp: every pixel written, counters consistent."""
    s = scenes["c3_masterchief"]
    P = s.render_params(width=1280, height=720, spp=8, seed=1)
    got, rgba, st = dev_scenes["c3_masterchief"].render(P)
    assert st.paths == 1280 * 720 * 8
    assert 1.0 <= st.segments / st.paths <= 4.0
    assert (got.sum(axis=2) > 0).mean() > 0.999
    assert rgba[0, 0, 2] > 200      # sky


def test_errors_are_loud(device, scenes):
    import ctypes as C
    from sexy_raytracer_b200 import abi
    from sexy_raytracer_b200.device import lib, SrtError
    s = scenes["c2_cube"]
    d = s.desc()
    bad = np.array(s.nodes)
    bad["left"][0] = 10 ** 6
    d.nodes = bad.ctypes.data_as(C.POINTER(abi.BvhNode))
    h = C.c_void_p()
    assert lib().srt_scene_create(device.h, C.byref(d), C.byref(h)) == 1
    assert b"out of range" in lib().srt_last_error()
    ds = device.upload(s)
    with pytest.raises(SrtError):
        ds.render(s.render_params(width=1))
    ds.close()


# ---- trees that stay in HBM: the compact 16-byte tree + leaf-ordered primitives (srt_core.cuh), and the 32-byte nodes ----
@pytest.fixture(scope="module")
def hbm_scenes(device, scenes):
    """The four reference scenes uploaded with their tree kept OUT of shared memory (SRT_FORCE_GLOBAL_TREE), which is
    what a tree beyond 227 KB gets (BASELINE config 5): srt_trace_rays and k_wf_extend then walk the compact tree."""
    import os
    os.environ["SRT_FORCE_GLOBAL_TREE"] = "1"
    try:
        out = {n: device.upload(s) for n, s in scenes.items()}
    finally:
        del os.environ["SRT_FORCE_GLOBAL_TREE"]
    yield out
    for s in out.values():
        s.close()


@pytest.mark.parametrize("name", SCENES)
def test_compact_tree_trace_vs_reference_golden(name, hbm_scenes, golden):
    """Same bar as the shared-memory path: ids bit-exact against the compiled reference on every golden ray set - the
    snapped-box filter may never change a decision."""
    g = golden.kat(name)
    for k in SETS:
        check_hits(hbm_scenes[name].trace(g["rays_" + k]), g["hits_" + k])
    check_hits(hbm_scenes[name].trace(g["rays_bounded"], 0.05, 0.9), g["hits_bounded"])


@pytest.mark.parametrize("name", SCENES)
def test_compact_tree_counters_and_fallback_rate(name, device, hbm_scenes, dev_scenes, golden):
    g = golden.kat(name)
    device.set_counters(True)
    try:
        hbm_scenes[name].trace(g["rays_secondary"])
        c = device.last_counters()
        dev_scenes[name].trace(g["rays_secondary"])
        c0 = device.last_counters()
    finally:
        device.set_counters(False)
    box, leaf = (int(x) for x in g["count_secondary"])
    assert c["node_visits"] == box == c0["node_visits"]
    assert c["leaf_tests"] == c0["leaf_tests"] and c["sphere_leaf_tests"] == c0["sphere_leaf_tests"]
    # the undecided shell of the snapped boxes: a minority of the tests even on these small scenes (the 1000-unit ground
    # sphere stretches the grid of the default scene to 0.03 units per cell)
    assert c["exact_box_tests"] < 0.35 * box


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("tree,sort", [("compact", 0), ("full", 0), ("compact", 1), ("full", 1)])
def test_hbm_tree_wavefront_agrees(name, tree, sort, hbm_scenes, dev_scenes, scenes):
    """k_wf_extend over the compact tree (MODE 2) and over the 32-byte nodes in HBM (MODE 1), with and without the rays of a
    wave sorted by origin cell and direction octant (k_wf_keys), trace the same paths as the megakernel over the
    shared-memory tree."""
    import os
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=11)
    a, ra, sa = dev_scenes[name].render(s.render_params(variant=1, **kw))
    os.environ["SRT_WF_TREE"] = tree
    os.environ["SRT_WF_SORT"] = str(sort)
    try:
        b, rb, sb = hbm_scenes[name].render(s.render_params(variant=2, **kw))
    finally:
        del os.environ["SRT_WF_TREE"], os.environ["SRT_WF_SORT"]
    assert sa.paths == sb.paths and sa.segments == sb.segments
    fin = np.isfinite(a).all(axis=2)
    assert (np.isfinite(b).all(axis=2) == fin).all()
    assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert (np.abs(ra.astype(int) - rb.astype(int)) <= 1).all()
