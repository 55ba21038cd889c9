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
