"""BASELINE config 5 family (synthetic M-triangle mesh + N spheres, tree from the bvh.h-equivalent builder) at sizes the
oracles finish in seconds.  CPU: the pinned C port against the COMPILED REFERENCE on such a tree.  GPU: the global-memory
(tree does not fit shared memory) code paths against the port."""
import numpy as np
import pytest

import rayset
from sexy_raytracer_b200.synth import make_stress_scene


@pytest.fixture(scope="module")
def small():
    return make_stress_scene(3000, 60, seed=5, width=160, height=90, spp=4, depth=16, extent=6.0, threads=2)


@pytest.fixture(scope="module")
def medium():            # 20 k + 204 nodes x 32 B = 0.9 MB > 227 KB: nodes stay in global memory
    return make_stress_scene(20000, 200, seed=6, width=192, height=108, spp=4, depth=16, extent=12.0, threads=2)


def camera_rays(s, n, seed):
    """rays from the camera position into the scene + rays from inside it"""
    rng = np.random.default_rng(seed)
    r = rayset.random_rays(n, seed)
    r["o"][: n // 2] = s.cam_args[0:3] * np.float32(0.2)
    r["dir"][: n // 2] = -r["o"][: n // 2] + rng.uniform(-6, 6, (n // 2, 3)).astype(np.float32)
    return r


def test_tree_shape_and_determinism(small):
    n = small.n_prim
    again = make_stress_scene(3000, 60, seed=5, width=160, height=90, spp=4, depth=16, extent=6.0, threads=5)
    assert again.nodes.tobytes() == small.nodes.tobytes()
    leaves = np.concatenate([~small.nodes["left"][small.nodes["left"] < 0],
                             ~small.nodes["right"][(small.nodes["right"] < 0) & (small.nodes["right"] != -2**31)]])
    assert sorted(leaves.tolist()) == list(range(n))          # every primitive exactly once


def test_port_equals_reference_on_built_tree(small, port, ref):
    hp, hr = port.from_scene(small), ref.from_scene(small)
    rays = camera_rays(small, 30000, 1)
    a, b = port.trace(hp, rays), ref.trace(hr, rays)
    assert (a["prim"] == b["prim"]).all() and (a["prim"] >= 0).mean() > 0.2
    hit = a["prim"] >= 0
    for f in ("t", "p", "normal", "uv"):
        assert a[f][hit].tobytes() == b[f][hit].tobytes(), f
    assert port.count_tests(hp, rays) == ref.count_tests(hr, rays)
    port.free(hp); ref.free(hr)


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["small", "medium"])
def test_gpu_trace_on_built_tree(which, small, medium, device, port):
    from test_gpu_parity import check_hits
    s = small if which == "small" else medium
    ds, h = device.upload(s), port.from_scene(s)
    rays = camera_rays(s, 60000, 2)
    want = port.trace(h, rays)
    check_hits(ds.trace(rays), want)
    sec = rayset.secondary(want, 40000, 3)
    check_hits(ds.trace(sec), port.trace(h, sec))
    port.free(h); ds.close()


@pytest.mark.gpu
def test_gpu_frame_variants_on_global_tree(medium, device, port):
    """Tree in global memory (SMEM=false instantiations): megakernel, phased and wavefront trace the same paths as the port."""
    s = medium
    ds, h = device.upload(s), port.from_scene(s)
    P = s.render_params(seed=3)
    want, st_w = port.render(h, s, P, nthreads=8)
    port.free(h)
    outs = {}
    for v in (1, 2, 3):
        got, _, st = ds.render(s.render_params(seed=3, variant=v), want_rgba=False)
        assert st.variant == v and st.paths == st_w.paths
        assert abs(int(st.segments) - int(st_w.segments)) <= 2e-3 * st_w.segments
        fin = np.isfinite(want).all(axis=2) & np.isfinite(got).all(axis=2)
        assert abs(got[fin].sum() - want[fin].sum()) <= 5e-3 * want[fin].sum()
        outs[v] = (got, st)
    assert outs[1][1].segments == outs[2][1].segments == outs[3][1].segments
    assert (outs[1][0].view(np.uint32) == outs[3][0].view(np.uint32)).all()
    ds.close()
