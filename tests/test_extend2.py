"""Round-2 parity tests through the C ABI (all need the GPU):

  * writeColorTarget (color.h:25-41): the reference's own bytes for the reference's own sums
  * aabb::hit (aabb.h:11-27) through every box predicate the device uses, on the reference-generated KAT and on a generated
    set aimed at the predicates' decision thresholds
  * the production EXTEND kernels of the wavefront on explicit rays (srt_trace_rays_wavefront): k_wf_extend2 with 16- and
    32-byte nodes, k_wf_extend, against the C port (pinned to the compiled reference by tests/test_oracle.py)
  * Russian roulette (north_star; absent in the reference, main.cpp:36-37): same stream as the port, and unbiased
  * BASELINE config 5 at FULL size (10 M triangles + 100 k spheres): ids bit-exact through srt_trace_rays and through one
    production wave (compact tree, sorted rays, 2 x 768 threads per SM)
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import rayset
from conftest import SCENES, GOLDEN

pytestmark = pytest.mark.gpu


class env:
    def __init__(self, **kw):
        self.kw = kw

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kw}
        for k, v in self.kw.items():
            os.environ[k] = str(v)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v


# ---- writeColorTarget -----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", SCENES)
def test_resolve_equals_reference_bytes(name, device, golden):
    """srt_film_load(the reference's sums) -> srt_film_resolve == the bytes writeColorTarget produced for them."""
    from sexy_raytracer_b200.device import Film
    g = golden.image(name)
    H, W, _ = g["sum_a"].shape
    f = Film(device, W, H)
    try:
        for which in ("a", "b"):
            f.load(g["sum_" + which], int(g["spp"]))
            _, rgba = f.resolve(want_sum=False)
            assert (rgba == g["rgba_" + which]).all(), "%d bytes differ" % (rgba != g["rgba_" + which]).sum()
    finally:
        f.close()
    g = np.load(os.path.join(GOLDEN, "c3_masterchief_1k_image.npz"))
    if name == "c3_masterchief":
        f = Film(device, 426, 240)
        try:
            f.load(g["sum_a"], 1000)
            _, rgba = f.resolve(want_sum=False)
            assert (rgba == g["rgba_a"]).all()
        finally:
            f.close()


# ---- aabb::hit --------------------------------------------------------------------------------------------------------
def check_predicates(device, rays, boxes, tmm, want):
    for pred in (0, 1):
        dec, path = device.aabb_kat(rays, boxes, tmm, pred)
        assert (dec == want).all(), "predicate %d: %d decisions differ from aabb::hit" % (pred, (dec != want).sum())
    fb = {}
    for pred in (2, 3):
        dec, path = device.aabb_kat(rays, boxes, tmm, pred)
        wrong = (dec == 0) & (want == 1)
        assert not wrong.any(), "conservative cull (predicate %d) rejected %d boxes the reference hits" % (pred, wrong.sum())
        served = path == 0
        fb[pred] = (served.mean(), (dec[served & (want == 0)] == 0).mean() if (served & (want == 0)).any() else 1.0)
    return fb


def test_aabb_predicates_vs_reference_golden(device, golden):
    g = golden.aabb()
    tmm = g["tminmax"]
    assert (tmm[:, 0] == np.float32(0.001)).all()
    fb = check_predicates(device, g["rays"], g["boxes"], tmm, g["hit"])
    # the conservative test must still cull: most of the reference's misses, for the rays it serves
    assert fb[2][1] > 0.85 and fb[3][1] > 0.85, fb
    # the filtered predicate decides nearly everything without the reference's divisions
    dec, path = device.aabb_kat(g["rays"], g["boxes"], tmm, 0)
    ok = (g["rays"]["dir"] != 0).all(axis=1)
    assert path[ok].mean() < 0.10


def test_aabb_predicates_at_their_thresholds(device, port):
    """rays grazing faces / edges / corners within a few ulp, tMax equal to the exit distance, |o/d| up to 2^59, zero and
    denormal direction components; reference decisions from the C port (equal to the compiled reference's on this very set:
    tests/test_oracle.py::test_port_aabb_adversarial_vs_reference)."""
    for seed in (11, 12):
        rays, boxes, tmm = rayset.aabb_adversarial(300000, seed)
        want = port.aabb_kat(rays, boxes, tmm)
        assert 0.2 < want.mean() < 0.6
        check_predicates(device, rays, boxes, tmm, want)


# ---- the production extend kernels on explicit rays ----------------------------------------------------------------------
def port_trace_mt(port, h, rays, threads=8):
    parts = np.array_split(np.arange(len(rays)), threads * 4)
    with ThreadPoolExecutor(threads) as ex:
        out = list(ex.map(lambda idx: port.trace(h, rays[idx]), parts))
    return np.concatenate(out)


def ray_sets(s, want_fn, n=60000):
    pin = rayset.pinhole(s)
    wp = want_fn(pin)
    return np.concatenate([pin, rayset.secondary(wp, n, 5), rayset.random_rays(n, 6), rayset.adversarial(s, n // 2, 7)])


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("x2", [16, 32, 0])
def test_wavefront_extend_on_explicit_rays(name, x2, device, dev_scenes, scenes, port):
    """ids BIT-EXACT and t within 1e-5 of the port for every extend kernel the wavefront can pick for a tree in shared memory;
    the conservative walk visits at least the reference's nodes and tests exactly the reference's primitives."""
    s = scenes[name]
    h = port.from_scene(s)
    rays = ray_sets(s, lambda r: port.trace(h, r))
    want = port_trace_mt(port, h, rays)
    box, leaf = port.count_tests(h, rays[:50000])
    port.free(h)
    with env(SRT_WF_X2=x2):
        prim, t = dev_scenes[name].trace_wavefront(rays)
        device.set_counters(True)
        try:
            dev_scenes[name].trace_wavefront(rays[:50000])
            c = device.last_counters()
        finally:
            device.set_counters(False)
    assert (prim == want["prim"]).all(), "ids differ at %s" % np.flatnonzero(prim != want["prim"])[:10]
    hit = want["prim"] >= 0
    assert (t[hit].view(np.uint32) == want["t"][hit].view(np.uint32)).all() or \
        (np.abs(t[hit] - want["t"][hit]) <= 1e-5 * np.abs(want["t"][hit])).all()
    if x2 == 0:
        assert c["node_visits"] == box
    else:
        # boxes rounded outward (to fp16 in the 16-byte nodes) let a few more nodes through; the walk tree leaves out inner
        # nodes that cull too little (srt_scene_create), so it usually looks at FEWER boxes than the reference - never at
        # other leaves (the leaf-test counters below)
        assert c["node_visits"] <= (1.08 if x2 == 16 else 1.02) * box + 64, (c["node_visits"], box)
    assert c["leaf_tests"] <= leaf          # span-1 nodes test their primitive once (srt.h: SRT_BVH_NONE)
    with env(SRT_WF_X2=0):
        device.set_counters(True)
        try:
            dev_scenes[name].trace_wavefront(rays[:50000])
            c0 = device.last_counters()
        finally:
            device.set_counters(False)
    assert c["leaf_tests"] == c0["leaf_tests"] and c["sphere_leaf_tests"] == c0["sphere_leaf_tests"]


@pytest.mark.parametrize("x2", [16, 32])
def test_wavefront_extend2_small_pools_and_threads(x2, dev_scenes, scenes, port):
    """few rays (warps that never fill their park lists), other CTA sizes and walk thresholds: same hits"""
    s = scenes["c3_masterchief"]
    h = port.from_scene(s)
    rays = ray_sets(s, lambda r: port.trace(h, r), 4000)
    want = port.trace(h, rays)
    port.free(h)
    for n in (1, 31, 33, 1000, len(rays)):
        for wf in ("1048576,1024,16", "1048576,512,28", "1048576,256,4", "1048576,768,32", "1048576,1024,1"):
            with env(SRT_WF_X2=x2, SRT_WF=wf):
                prim, t = dev_scenes["c3_masterchief"].trace_wavefront(rays[:n])
            assert (prim == want["prim"][:n]).all(), (n, wf)


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("x2", [16, 32])
def test_wavefront_frame_with_extend2(name, x2, device, dev_scenes, scenes):
    """whole frames: variant 2 over k_wf_extend2 traces the same paths as the flat megakernel (same Philox keys)"""
    from test_gpu_parity import SMALL
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=11)
    a, ra, sa = dev_scenes[name].render(s.render_params(variant=1, **kw))
    with env(SRT_WF_X2=x2):
        b, rb, sb = dev_scenes[name].render(s.render_params(variant=2, **kw))
        with env(SRT_WF="4096,256,12"):
            c, _, sc_ = dev_scenes[name].render(s.render_params(variant=2, **kw), want_rgba=False)
    assert sa.paths == sb.paths and sa.segments == sb.segments == sc_.segments
    fin = np.isfinite(a).all(axis=2)
    assert (np.isfinite(b).all(axis=2) == fin).all()
    assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert np.abs(a[fin] - c[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert (np.abs(ra.astype(int) - rb.astype(int)) <= 1).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("x2", [16, 32])
def test_fused_path_kernel_frames(name, x2, device, dev_scenes, scenes):
    """k_wf_path (extend + shade + regenerate in ONE persistent launch, srt_path.cuh) traces the same paths as the flat
    megakernel: same path and segment counts, per-pixel sums equal up to fp32 summation order, RGBA8 within 1."""
    from test_gpu_parity import SMALL
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=11)
    a, ra, sa = dev_scenes[name].render(s.render_params(variant=1, **kw))
    with env(SRT_WF_X2=x2, SRT_WF_FUSED=1):
        b, rb, sb = dev_scenes[name].render(s.render_params(variant=2, **kw))
        with env(SRT_WF="4096,512,12"):
            c, _, sc_ = dev_scenes[name].render(s.render_params(variant=2, **kw), want_rgba=False)
    if sb.extend_kernel:                       # (a tree k_wf_extend2 cannot serve keeps the two-kernel wavefront)
        assert sb.fused == 1 and sb.launches == 2 and sc_.fused == 1      # k_wf_path + k_resolve
    assert sa.paths == sb.paths == sc_.paths and sa.segments == sb.segments == sc_.segments
    fin = np.isfinite(a).all(axis=2)
    assert (np.isfinite(b).all(axis=2) == fin).all()
    assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert np.abs(a[fin] - c[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max())
    assert (np.abs(ra.astype(int) - rb.astype(int)) <= 1).all()


@pytest.mark.gpu
def test_fused_path_kernel_odd_sizes_depths_and_counters(device, dev_scenes, scenes):
    """frames whose width / height are not multiples of the 8x4 tile, one-bounce and deep paths, Russian roulette, a sample
    window, and the counted build: k_wf_path against the two-kernel wavefront on the same Philox keys"""
    s = scenes["c3_masterchief"]
    ds = dev_scenes["c3_masterchief"]
    cases = [dict(width=37, height=23, spp=9, max_bounce=4), dict(width=130, height=66, spp=3, max_bounce=1),
             dict(width=64, height=36, spp=16, max_bounce=12, rr_begin=2), dict(width=2, height=2, spp=1, max_bounce=4),
             dict(width=100, height=56, spp=20, max_bounce=4, sample_begin=7, sample_count=5)]
    for kw in cases:
        with env(SRT_WF_FUSED=0):
            a, _, sa = ds.render(s.render_params(variant=2, seed=5, **kw), want_rgba=False)
        with env(SRT_WF_FUSED=1):
            b, _, sb = ds.render(s.render_params(variant=2, seed=5, **kw), want_rgba=False)
        assert sb.fused == 1 and sa.fused == 0, kw
        assert sa.paths == sb.paths and sa.segments == sb.segments, kw
        fin = np.isfinite(a).all(axis=2)
        assert (np.isfinite(b).all(axis=2) == fin).all()
        assert np.abs(a[fin] - b[fin]).max() <= 2e-5 * max(1.0, np.abs(a[fin]).max()), kw
    device.set_counters(True)
    try:
        kw = dict(width=160, height=90, spp=8, max_bounce=4, seed=5)
        with env(SRT_WF_FUSED=0):
            _, _, sa = ds.render(s.render_params(variant=2, **kw), want_rgba=False, want_sum=False)
            ca = device.last_counters()
        with env(SRT_WF_FUSED=1):
            _, _, sb = ds.render(s.render_params(variant=2, **kw), want_rgba=False, want_sum=False)
            cb = device.last_counters()
    finally:
        device.set_counters(False)
    assert sa.segments == sb.segments
    for k in ("leaf_tests", "sphere_leaf_tests", "node_visits", "exact_box_tests", "triangle_hits"):
        assert ca[k] == cb[k], (k, ca[k], cb[k])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1_shirley", "c3_masterchief"])
def test_shade_kernels_agree(name, dev_scenes, scenes):
    """k_wf_shade (one thread per slot), k_wf_shade2 (sorted passes per 512-slot chunk) and k_wf_shade2c (the default: lists that
    carry their remainder to the next chunk) trace the same paths: same path and segment counts, sums equal up to fp32 order.
    Small pools make blocks serve several chunks, so the carried lists are exercised."""
    from test_gpu_parity import SMALL
    s = scenes[name]
    W, H, spp = SMALL[name]
    kw = dict(width=W, height=H, spp=spp, seed=23, variant=2)
    ref = None
    for kind, wf in ((1, None), (2, None), (3, None), (3, "4096,256,8"), (3, "65536,1024,8"), (2, "4096,512,8")):
        e = dict(SRT_WF_SHADE=kind)
        if wf:
            e["SRT_WF"] = wf
        with env(**e):
            a, _, st = dev_scenes[name].render(s.render_params(**kw), want_rgba=False)
        if ref is None:
            ref = (a, st)
            continue
        assert st.paths == ref[1].paths and st.segments == ref[1].segments, (kind, wf)
        fin = np.isfinite(ref[0]).all(axis=2)
        assert (np.isfinite(a).all(axis=2) == fin).all()
        assert np.abs(a[fin] - ref[0][fin]).max() <= 2e-5 * max(1.0, np.abs(ref[0][fin]).max()), (kind, wf)


@pytest.mark.gpu
def test_walk_tree(device, scenes, port):
    """Scene::x2_nodes, the tree k_wf_extend2 walks.  With the reference's own inner nodes, all kept (SRT_X2_SAH=0, SRT_X2_TAU=9)
    the conservative walk looks at the reference's boxes or a few more.  The default - a surface-area sweep over the reference's
    leaf sequence, then inner nodes with an area ratio above 0.6 left out - looks at about half as many, and at exactly the
    same leaves and primitives: same ids, same leaf-test counters."""
    s = scenes["c3_masterchief"]
    h = port.from_scene(s)
    try:
        rays = ray_sets(s, lambda r: port.trace(h, r), 20000)
        want = port.trace(h, rays)
        box, leaf = port.count_tests(h, rays)
    finally:
        port.free(h)
    out = {}
    for key, sah, tau in (("ref", 0, "9"), ("ref+collapse", 0, "0.6"), ("sah", 1, "9"), ("default", 1, "0.6"), ("sah 0.3", 1, "0.3")):
        with env(SRT_X2_TAU=tau, SRT_X2_SAH=sah):
            ds = device.upload(s)
        try:
            device.set_counters(True)
            for x2 in (32, 16):
                with env(SRT_WF_X2=x2):
                    prim, t = ds.trace_wavefront(rays)
                assert (prim == want["prim"]).all(), (key, x2)
            with env(SRT_WF_X2=32):
                prim, t = ds.trace_wavefront(rays)
            out[key] = device.last_counters()
        finally:
            device.set_counters(False)
            ds.close()
    for key in out:
        assert out[key]["leaf_tests"] == out["ref"]["leaf_tests"] and out[key]["sphere_leaf_tests"] == out["ref"]["sphere_leaf_tests"], key
    assert box <= out["ref"]["node_visits"] <= 1.02 * box + 64
    assert out["ref+collapse"]["node_visits"] < 0.85 * box
    assert out["default"]["node_visits"] < 0.65 * box, (out["default"]["node_visits"], box)
    msg = "box tests per ray: reference %.1f; walk tree %s" % (box / len(rays), ", ".join("%s %.1f" % (k, out[k]["node_visits"] / len(rays)) for k in out))
    print(msg)
    path = os.environ.get("SRT_TEST_REPORT")
    if path:
        with open(path, "a") as f:
            f.write(msg + "\n")


def test_extend2_refuses_trees_it_cannot_serve(device, scenes, port):
    """a tree whose boxes are not nested keeps k_wf_extend (plan_wavefront), with the same hits"""
    import copy
    s = copy.copy(scenes["c2_cube"])
    s.nodes = np.array(s.nodes)
    inner = np.flatnonzero(s.nodes["left"] >= 0)
    c = int(s.nodes["left"][inner[0]])
    s.nodes["bmax"][c] += np.float32(0.5)            # the child now sticks out of its parent: legal for hittable, not nested
    h = port.from_scene(s)
    ds = device.upload(s)
    try:
        rays = ray_sets(s, lambda r: port.trace(h, r), 20000)
        want = port.trace(h, rays)
        device.set_counters(True)
        try:
            prim, t = ds.trace_wavefront(rays)
            c1 = device.last_counters()
        finally:
            device.set_counters(False)
        box, leaf = port.count_tests(h, rays)
        assert (prim == want["prim"]).all()
        assert c1["node_visits"] == box                  # the exact walk ran, not the conservative one
    finally:
        port.free(h)
        ds.close()


# ---- Russian roulette ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rr", [1, 3])
@pytest.mark.parametrize("variant", [1, 2])
def test_russian_roulette_vs_port_same_stream(rr, variant, dev_scenes, scenes, port):
    s = scenes["c1_shirley"]
    P = s.render_params(width=200, height=112, spp=8, seed=5, rr_begin=rr, variant=variant)
    h = port.from_scene(s)
    want, st_w = port.render(h, s, P, nthreads=8)
    P0 = s.render_params(width=200, height=112, spp=8, seed=5, variant=variant)
    _, st_0 = port.render(h, s, P0, nthreads=8)
    port.free(h)
    got, _, st = dev_scenes["c1_shirley"].render(P, want_rgba=False)
    assert st.paths == st_w.paths
    assert abs(int(st.segments) - int(st_w.segments)) <= 1e-3 * st_w.segments
    assert st_w.segments < 0.97 * st_0.segments              # roulette does end paths early
    fin = np.isfinite(want).all(axis=2) & np.isfinite(got).all(axis=2)
    err = np.abs(got - want).max(axis=2) / np.maximum(np.abs(want).max(axis=2), 1e-3)
    assert (err[fin] > 1e-3).mean() < 0.01
    assert abs(got[fin].sum() - want[fin].sum()) <= 2e-3 * want[fin].sum()


def test_russian_roulette_is_unbiased(dev_scenes, scenes):
    """tile means with rr_begin = 2 against rr_begin = 0, 256 spp, two seeds each: the roulette frames deviate from the plain
    ones no more than plain frames with different seeds deviate from each other (x 1.5: roulette adds variance, not bias)."""
    s = scenes["c1_shirley"]
    ds = dev_scenes["c1_shirley"]
    W, H, spp = 400, 225, 256

    def tiles(seed, rr):
        acc, _, st = ds.render(s.render_params(width=W, height=H, spp=spp, seed=seed, rr_begin=rr), want_rgba=False)
        x = np.where(np.isfinite(acc), acc, 0)[: H // 16 * 16, : W // 16 * 16].astype(np.float64)
        return x.reshape(H // 16, 16, W // 16, 16, 3).mean(axis=(1, 3)) / spp, st
    a0, s0 = tiles(1, 0)
    b0, _ = tiles(2, 0)
    a2, s2 = tiles(3, 2)
    b2, _ = tiles(4, 2)
    assert s2.segments < 0.9 * s0.segments
    ref = 0.5 * (a0 + b0)
    n = np.abs(ref) + 1e-3
    spread = np.sqrt((((a0 - b0) / n) ** 2).mean())
    dev = np.sqrt((((0.5 * (a2 + b2) - ref) / n) ** 2).mean())
    assert dev <= 1.5 * spread, (dev, spread)
    assert abs((a2 + b2).sum() - (a0 + b0).sum()) <= 0.01 * (a0 + b0).sum()


def test_depth_zero_is_black(dev_scenes, scenes):
    """maxBounce <= 0: rayColor returns (0,0,0) without calling world.hit (main.cpp:36-37)"""
    s = scenes["c2_cube"]
    for v in (0, 1, 2, 3):
        acc, rgba, st = dev_scenes["c2_cube"].render(s.render_params(width=64, height=36, spp=3, max_bounce=0, variant=v))
        assert st.segments == 0 and st.paths == 64 * 36 * 3
        assert (acc == 0).all() and (rgba[..., :3] == 0).all() and (rgba[..., 3] == 255).all()


# ---- BASELINE config 5 at full size -----------------------------------------------------------------------------------
def test_config5_full_size_ids(device, port):
    """10 M triangles + 100 k spheres (SURVEY 8.4's generator, tree from the bvh.h-equivalent builder, 11.8 M nodes): primary,
    secondary and axis-parallel rays through srt_trace_rays (compact tree, exact fallback) and through ONE production wave
    of the wavefront (k_wf_extend<.,2,768>, rays sorted by octant + Morton cell), ids bit-exact and t within 1e-5 of the port."""
    from sexy_raytracer_b200.synth import make_stress_scene
    s = make_stress_scene()
    assert s.n_tri == 10_000_000 and s.n_sph == 100_016 and len(s.nodes) > 11_000_000
    ds = device.upload(s)
    h = port.from_scene(s)
    try:
        pin = rayset.pinhole(s, step=6)                    # 320 x 180 = 57 600 primary rays
        wp = port_trace_mt(port, h, pin)
        assert (wp["prim"] >= 0).mean() > 0.5
        rng = np.random.default_rng(8)
        sec = rayset.secondary(wp, 120000, 9)
        axis = rayset.secondary(wp, 24000, 10)
        z = rng.integers(0, 3, len(axis))
        axis["dir"][np.arange(len(axis)), z] = 0
        inside = rayset.random_rays(20000, 11, lo=-50.0, hi=50.0)
        rays = np.concatenate([pin, sec, axis, inside])
        assert len(rays) >= 200000
        want = port_trace_mt(port, h, rays)
        got = ds.trace(rays)
        assert (got["prim"] == want["prim"]).all(), "srt_trace_rays ids differ at %s" % np.flatnonzero(got["prim"] != want["prim"])[:10]
        hit = want["prim"] >= 0
        assert (np.abs(got["t"][hit] - want["t"][hit]) <= 1e-5 * np.abs(want["t"][hit])).all()
        # the undecided shell of the snapped boxes (0.0015-unit cells): rays the filter serves re-evaluate aabb::hit on < 5 % of
        # their box tests (axis-parallel rays always run the reference's arithmetic and are left out of this count)
        served = np.concatenate([pin, sec, inside])
        device.set_counters(True)
        try:
            ds.trace(served)
            c = device.last_counters()
        finally:
            device.set_counters(False)
        assert c["exact_box_tests"] < 0.05 * c["node_visits"], "exact fallback rate %.4f" % (c["exact_box_tests"] / c["node_visits"])
        prim, t = ds.trace_wavefront(rays)
        assert (prim == want["prim"]).all(), "wavefront ids differ at %s" % np.flatnonzero(prim != want["prim"])[:10]
        assert (np.abs(t[hit] - want["t"][hit]) <= 1e-5 * np.abs(want["t"][hit])).all()
        print("config 5 full size: %d rays, %.1f %% hit, %.0f box tests / ray, exact fallback %.2f %%" %
              (len(rays), 100 * hit.mean(), c["node_visits"] / len(rays), 100 * c["exact_box_tests"] / c["node_visits"]))
    finally:
        port.free(h)
        ds.close()
