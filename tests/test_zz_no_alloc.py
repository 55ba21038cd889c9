"""srt_trace_rays_wavefront_static: the wave of srt_trace_rays_wavefront for callers that may not allocate (include/srt.h; no
counterpart in the reference).  Collected LAST on purpose;
passes alone in a fresh process and after the other tests.

Called through the C ABI on device pointers (torch owns the caller's three buffers; the entry point itself allocates nothing) and
checked against the existing entry point on the same rays: primitive ids and t bit for bit (the tolerance of the parity tests for
integer work and for t of the same kernels), on two calls in a row that share the static scratch, the second with other and
fewer rays, for a tree in shared memory (k_wf_extend2) and for trees kept in HBM (k_wf_extend, both node formats)."""
import os

import numpy as np
import pytest

import rayset
from sexy_raytracer_b200 import abi
from sexy_raytracer_b200.device import SrtError, lib


def _rays(scene, dscene, n, seed):
    pin = rayset.pinhole(scene, step=2)
    hits = dscene.trace(pin)
    parts = [pin, rayset.secondary(hits, n // 2, seed), rayset.random_rays(n // 4, seed + 1), rayset.adversarial(scene, n // 4, seed + 2)]
    r = np.concatenate(parts)
    sel = np.random.RandomState(seed).permutation(len(r))[:n]
    return np.ascontiguousarray(r[sel], abi.RAY_DTYPE)


def _static_call(dscene, rays):
    import torch
    n = len(rays)
    d_rays = torch.from_numpy(rays.view(np.uint8).copy()).cuda()
    d_prim = torch.full((max(n, 1),), -77, dtype=torch.int32, device="cuda")
    d_t = torch.full((max(n, 1),), -1.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()                     # the library launches on its own stream
    dscene.trace_wavefront_static(d_rays.data_ptr(), n, d_prim.data_ptr(), d_t.data_ptr())
    return d_prim.cpu().numpy()[:n], d_t.cpu().numpy()[:n]


def _check(dscene, rays):
    want_prim, want_t = dscene.trace_wavefront(rays)
    got_prim, got_t = _static_call(dscene, rays)
    assert (got_prim == want_prim).all(), "ids differ at %s" % np.flatnonzero(got_prim != want_prim)[:10]
    assert (got_t.view(np.uint32) == want_t.view(np.uint32)).all()
    return want_prim


@pytest.mark.gpu
def test_static_entry_point_equals_the_allocating_one(device, scenes, dev_scenes):
    limit = int(lib().srt_static_max_rays())
    assert limit == 65536
    s = scenes["c3_masterchief"]
    ds = dev_scenes["c3_masterchief"]
    a = _rays(s, ds, 60000, 41)
    b = _rays(s, ds, 12345, 97)                  # other rays, fewer: the scratch still holds call 1's slots beyond them
    pa = _check(ds, a)
    pb = _check(ds, b)
    assert (pa >= 0).any() and (pa < 0).any() and (pb >= 0).any()
    _check(ds, a[:limit])                        # and again the first set: results do not depend on what the scratch held
    _check(ds, a[:1])
    # trees kept out of shared memory: k_wf_extend over the compact tree / the 32-byte nodes (never sorted here)
    os.environ["SRT_FORCE_GLOBAL_TREE"] = "1"
    try:
        hbm = device.upload(s)
    finally:
        del os.environ["SRT_FORCE_GLOBAL_TREE"]
    try:
        for tree in ("compact", "full"):
            os.environ["SRT_WF_TREE"] = tree
            try:
                _check(hbm, a)
                _check(hbm, b)
            finally:
                del os.environ["SRT_WF_TREE"]
    finally:
        hbm.close()
    # the small scenes (config 2's 15-node tree, spheres only)
    for name in ("c2_cube", "c1_shirley"):
        r = _rays(scenes[name], dev_scenes[name], 20000, 5)
        _check(dev_scenes[name], r)


@pytest.mark.gpu
def test_static_entry_point_limits(scenes, dev_scenes):
    import torch
    limit = int(lib().srt_static_max_rays())
    ds = dev_scenes["c3_masterchief"]
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(SrtError, match="SRT_STATIC_MAX_RAYS"):
        ds.trace_wavefront_static(buf.data_ptr(), limit + 1, buf.data_ptr(), buf.data_ptr())
    with pytest.raises(SrtError):
        ds.trace_wavefront_static(None, 5, buf.data_ptr(), buf.data_ptr())
    ds.trace_wavefront_static(None, 0, None, None)          # nothing to do
    # a refused call leaves the entry point usable
    r = _rays(scenes["c3_masterchief"], ds, 3000, 11)
    _check(ds, r)
