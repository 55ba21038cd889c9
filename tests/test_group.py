"""Several GPUs behind one handle (include/srt.h: srt_device_group / srt_scene_group / srt_render_group).

CPU: the sample split of the C library equals sexy_raytracer_b200.dist.sample_split (the multi-process route).
GPU: the group's frame equals the single-device frame.  On a single-GPU box the group is {0, 0, 0} - three srt_device
objects sharing GPU 0, which exercises the split, the per-device renders on their own streams and host threads, the peer
copies and the sum; with two or more GPUs it is the real thing."""
import ctypes as C

import numpy as np
import pytest

from sexy_raytracer_b200 import abi, build
from sexy_raytracer_b200.dist import sample_split


def test_c_sample_split_equals_python():
    build.build_cuda()
    L = abi.bind(C.CDLL(build.LIB))
    b, c = C.c_int32(), C.c_int32()
    for spp in (0, 1, 7, 16, 255, 5000):
        for n in (1, 2, 3, 4, 8):
            for k in range(n):
                assert L.srt_sample_split(0, spp, n, k, C.byref(b), C.byref(c)) == 0
                assert (b.value, c.value) == sample_split(spp, n, k)
    assert L.srt_sample_split(100, 10, 4, 3, C.byref(b), C.byref(c)) == 0 and (b.value, c.value) == (108, 2)
    assert L.srt_sample_split(0, 10, 4, 4, C.byref(b), C.byref(c)) != 0


def _gpu_ids():
    import torch
    n = torch.cuda.device_count()
    return list(range(n)) if n >= 2 else [0, 0, 0]


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [1, 2])
def test_group_frame_equals_single_device_frame(variant, dev_scenes, scenes):
    from sexy_raytracer_b200.device import DeviceGroup
    s = scenes["c3_masterchief"]
    ids = _gpu_ids()
    g = DeviceGroup(ids)
    sg = g.upload(s)
    try:
        assert len(g) == len(ids)
        for spp in (16, 7, 2):                      # 2 samples on 3 devices: one of them renders nothing
            P = s.render_params(width=320, height=180, spp=spp, seed=4, variant=variant)
            one, rgba1, st1 = dev_scenes["c3_masterchief"].render(P)
            acc, rgba, st = sg.render(P)
            assert st.paths == st1.paths == 320 * 180 * spp and st.segments == st1.segments
            fin = np.isfinite(one).all(axis=2)
            assert (np.isfinite(acc).all(axis=2) == fin).all()
            assert np.abs(acc[fin] - one[fin]).max() <= 1e-4 * max(1.0, np.abs(one[fin]).max())
            assert (np.abs(rgba.astype(int) - rgba1.astype(int)) <= 1).all() and (rgba[..., 3] == 255).all()
        # a sample sub-range through the group (progressive callers)
        P = s.render_params(width=320, height=180, spp=16, sample_begin=5, sample_count=9, seed=4, variant=variant)
        one, _, st1 = dev_scenes["c3_masterchief"].render(P, want_rgba=False)
        acc, _, st = sg.render(P, want_rgba=False)
        assert st.segments == st1.segments
        fin = np.isfinite(one).all(axis=2)
        assert np.abs(acc[fin] - one[fin]).max() <= 1e-4 * max(1.0, np.abs(one[fin]).max())
    finally:
        sg.close()
        g.close()


@pytest.mark.gpu
def test_group_default_takes_every_gpu_and_errors_are_loud():
    import torch
    from sexy_raytracer_b200.device import DeviceGroup, SrtError
    g = DeviceGroup()
    assert len(g) == torch.cuda.device_count()
    g.close()
    with pytest.raises(SrtError):
        DeviceGroup([torch.cuda.device_count() + 3])


@pytest.mark.gpu
def test_example_uses_the_group(tmp_path):
    """examples/sexy-raytracer (b200Device::init's default = every visible GPU) renders through srt_render_group"""
    import os
    import subprocess
    import torch
    exe = build.build_example()
    env = dict(os.environ, SRT_B200_LIBRARY=build.LIB)
    r = subprocess.run([exe, "shirley", "90", "8", str(tmp_path / "g.png")], capture_output=True, text=True, env=env)
    assert r.returncode == 0 and ("rendering on %d GPU(s)" % torch.cuda.device_count()) in r.stderr
