"""Host-side helpers (libsrt_host.so) against what the COMPILED REFERENCE produced: the committed scene files carry the
tree its bvhNode constructor built and the camera its constructor derived (oracle/make_golden.py -> ref_scene_export)."""
import numpy as np
import pytest

from conftest import SCENES
from sexy_raytracer_b200 import host


def same_nodes(a, b):
    return a.shape == b.shape and a.tobytes() == b.tobytes()


@pytest.mark.parametrize("name", SCENES)
def test_builder_reproduces_reference_tree(name, scenes):
    """Same RNG stream, comparator and std::sort => the identical tree, bit for bit (boxes, structure, leaf ids).
    The number of randomFloat() draws the reference made BEFORE the constructor (scene generation) is found by search."""
    s = scenes[name]
    # 4431 = the draws of the Shirley block (main.cpp:92-122) for the 485-sphere scene, found by this search once
    for skip in [0, 4431] + list(range(1, 6000)):
        nodes, draws = host.build_bvh(s, 0.0, 1.0, rng_skip=skip, threads=2)
        assert draws == len(s.nodes) == len(nodes)
        if same_nodes(nodes, np.ascontiguousarray(s.nodes)):
            break
    else:
        pytest.fail("no RNG offset below 6000 reproduces the reference tree of %s" % name)
    assert skip == (4431 if name == "c1_shirley" else 0)
    if False:
        assert skip == 0          # main.cpp's default branch draws nothing before bvhNode (main.cpp:146)
    # thread count must not matter
    again, _ = host.build_bvh(s, 0.0, 1.0, rng_skip=skip, threads=7)
    assert same_nodes(again, nodes)


@pytest.mark.parametrize("name", SCENES)
def test_camera_constructor_bit_exact(name, scenes):
    s = scenes[name]
    a = s.cam_args
    cam = host.make_camera(a[0:3], a[3:6], a[6:9], *[float(x) for x in a[9:15]])
    assert cam.tobytes() == np.ascontiguousarray(s.camera, np.float32).tobytes()


def test_builder_edge_cases(scenes):
    from sexy_raytracer_b200.scene import Scene
    e = Scene()
    nodes, draws = host.build_bvh(e)
    assert len(nodes) == 0 and draws == 0
    s = scenes["c2_cube"]
    one = Scene()
    for k in ("sph_c0", "sph_c1", "sph_t0", "sph_t1", "sph_r", "sph_mat"):
        setattr(one, k, getattr(s, k)[:1])
    nodes, draws = host.build_bvh(one)
    assert len(nodes) == 1 and nodes["left"][0] == ~0 and nodes["right"][0] == -2**31
    boxes = host.prim_boxes(one)
    assert (nodes["bmin"][0] == boxes[0, :3]).all() and (nodes["bmax"][0] == boxes[0, 3:]).all()
