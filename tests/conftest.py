import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from sexy_raytracer_b200.scene import Scene  # noqa: E402

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
