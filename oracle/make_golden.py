"""Generates scenes/*.npz and tests/golden/*.npz FROM THE REFERENCE (oracle/_ref/libref_oracle.so, i.e.
/root/reference's own code compiled by oracle/Makefile).  Run here, in the container that has
/root/reference; the outputs are committed so that the GPU box (no /root/reference) can check against them.

    python oracle/make_golden.py            # everything (about 3 minutes of reference CPU rendering)
    python oracle/make_golden.py --no-images

Scenes: c1 Shirley spheres, c2 cube+square glTF, c3 main.cpp default scene (Master Chief +
globe; the four globe maps are missing from this checkout -> the reference's (1,0,1) fallback),
c3s = c3 with the globe slots pointed at files that exist (1-byte-per-texel and normal-map paths).
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle.refbind import RefOracle  # noqa: E402
import rayset  # noqa: E402

F = np.float32
BG = np.array([0.53, 0.81, 0.92], F)     # main.cpp:170

# name: (native id, cam args [eye3 lookAt3 up3 vfov aspect aperture focus t0 t1], (W, H, spp, depth))
CONFIGS = {
    "c1_shirley": (1, [13, 2, 3, 0, 0, 0, 0, 1, 0, 20.0, 16.0 / 9.0, 0.1, 10.0, 0, 1.0], (400, 225, 16, 10)),
    "c2_cube": (2, [3, 3, 5, 0, 0.5, 0, 0, 1, 0, 40.0, 16.0 / 9.0, 0.0, 10.0, 0, 1.0], (640, 360, 64, 8)),
    "c3_masterchief": (3, [0, 3, 5, 0, 2.5, 0, 0, 1, 0, 70.0, 16.0 / 9.0, 0.1, 10.0, 0, 1.0], (426, 240, 1000, 4)),
    "c3s_globe": (4, [0, 3, 5, 0, 2.5, 0, 0, 1, 0, 70.0, 16.0 / 9.0, 0.1, 10.0, 0, 1.0], (426, 240, 1000, 4)),
}
# golden reference renders: (W, H, spp) small enough for the CPU
IMAGES = {"c1_shirley": (400, 225, 16), "c2_cube": (320, 180, 32), "c3_masterchief": (426, 240, 32),
          "c3s_globe": (213, 120, 32)}


def grid24(rng, shape):
    """uniform numbers on the 2^-24 grid (what randomFloat() can return)"""
    return (rng.integers(0, 1 << 24, shape).astype(np.float64) / float(1 << 24)).astype(F)


def ball_xi(rng, n):
    """xi triples whose image -1+2*xi lies strictly inside the unit ball (first rejection iteration accepts)"""
    out = np.zeros((n, 3), F)
    k = 0
    while k < n:
        xi = grid24(rng, (2 * n, 3))
        p = (F(-1) + F(2) * xi).astype(F)
        ls = ((p[:, 0] * p[:, 0] + p[:, 1] * p[:, 1]).astype(F) + p[:, 2] * p[:, 2]).astype(F)
        xi = xi[ls < F(0.98)]
        m = min(n - k, len(xi))
        out[k:k + m] = xi[:m]
        k += m
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-images", action="store_true")
    ap.add_argument("--aabb-only", action="store_true")
    ap.add_argument("--sampler-only", action="store_true")
    args = ap.parse_args()
    R = RefOracle()
    if args.sampler_only:
        # main.cpp:205-207 as the reference computes it, samples 0..63 (the angle is not reduced mod 360)
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sampler_kat.npz"),
                            pos=np.stack([R.rotated_grid(s) for s in range(64)]))
        return
    os.makedirs(os.path.join(ROOT, "scenes"), exist_ok=True)
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    rng = np.random.default_rng(20261104)

    for name, (which, cam_args, render) in ({} if args.aabb_only else CONFIGS).items():
        t0 = time.time()
        h = R.native(which)
        s = R.export(h)
        s.name = name
        s.cam_args = np.array(cam_args, F)
        s.camera = R.camera(s.cam_args)
        s.render = np.array(render, np.int32)
        s.background = BG.copy()
        s.save(os.path.join(ROOT, "scenes", name + ".npz"))

        # ---- hit records
        pin = rayset.pinhole(s, step=max(1, int(np.sqrt(render[0] * render[1] / 5000.0))))
        hp = R.trace(h, pin)
        sets = {"pinhole": pin, "secondary": rayset.secondary(hp, 4000, 11), "random": rayset.random_rays(2500, 12),
                "adversarial": rayset.adversarial(s, 2400, 13)}
        out = {}
        for k, rays in sets.items():
            out["rays_" + k] = rays
            out["hits_" + k] = R.trace(h, rays)
        # a bounded interval too (hittableList::hit passes tMax through)
        out["rays_bounded"] = sets["random"]
        out["hits_bounded"] = R.trace(h, sets["random"], 0.05, 0.9)
        box, leaf = R.count_tests(h, sets["secondary"])
        out["count_secondary"] = np.array([box, leaf], np.int64)

        # ---- material::scatter / emitted with injected random numbers
        allr = np.concatenate([sets["pinhole"], sets["secondary"], sets["random"]])
        allh = np.concatenate([out["hits_pinhole"], out["hits_secondary"], out["hits_random"]])
        ok = np.flatnonzero((allh["prim"] >= 0) & np.isfinite(allh["uv"]).all(axis=1))
        pick = rng.choice(ok, size=min(4000, len(ok)), replace=False)
        xi = np.concatenate([ball_xi(rng, len(pick)), grid24(rng, (len(pick), 1))], axis=1)
        att, emit, sc_rays, sc = R.shade_kat(h, allr[pick], allh[pick], xi)
        out.update(shade_rays=allr[pick], shade_hits=allh[pick], shade_xi=xi, shade_att=att, shade_emit=emit,
                   shade_out=sc_rays, shade_scattered=sc)

        # ---- texture::value
        for t in range(len(s.tex_kind)):
            if name == "c1_shirley" and t > 8:
                break
            uvp = np.concatenate([rng.uniform(-0.1, 1.1, (256, 2)), rng.uniform(-5, 5, (256, 3))], axis=1).astype(F)
            uvp[:8, :2] = [[0, 0], [1, 1], [0, 1], [1, 0], [0.5, 0.5], [0.999999, 0.000001], [1.0, 0.5], [0.5, 1.0]]
            out["tex_uvp_%d" % t] = uvp
            out["tex_val_%d" % t] = R.texture_kat(h, t, uvp)

        # ---- camera::getRay
        gin = np.zeros((1000, 5), F)
        gin[:, 0:2] = rng.uniform(-0.05, 1.05, (1000, 2))
        d = ball_xi(rng, 1000)[:, :2]                    # inside the unit disk as well
        gin[:, 2:4] = d
        gin[:, 4] = grid24(rng, 1000)
        gr = R.getray_kat(s.cam_args, gin)
        disk = (F(-1) + F(2) * gin[:, 2:4]).astype(F)    # randomFloat(-1,1) = -1 + 2*xi
        gk = gin.copy()
        gk[:, 2:4] = disk
        out.update(getray_in=gk, getray_out=gr)

        np.savez_compressed(os.path.join(gold, name + "_kat.npz"), **out)

        # ---- reference frames, two independent mt19937 streams (A, B)
        if not args.no_images:
            W, H, spp = IMAGES[name]
            a, ra, seg_a = R.render(h, s.cam_args, W, H, spp, render[3], BG, seed=5489)
            b, rb, seg_b = R.render(h, s.cam_args, W, H, spp, render[3], BG, seed=20261104)
            np.savez_compressed(os.path.join(gold, name + "_image.npz"), sum_a=a, sum_b=b, rgba_a=ra, rgba_b=rb,
                                spp=np.int32(spp), depth=np.int32(render[3]), segments=np.array([seg_a, seg_b], np.int64))
        R.free(h)
        print(name, "tri", s.n_tri, "sph", s.n_sph, "nodes", len(s.nodes), "%.1fs" % (time.time() - t0), flush=True)

    # ---- aabb::hit on explicit boxes (incl. zero direction components, degenerate and touching boxes)
    n = 20000
    o = rng.uniform(-3, 3, (n, 3)).astype(F)
    lo = rng.uniform(-2, 2, (n, 3)).astype(F)
    hi = (lo + rng.uniform(0, 2, (n, 3)).astype(F) * (rng.random((n, 3)) > 0.05)).astype(F)
    aim = (lo + (hi - lo) * rng.uniform(-0.15, 1.15, (n, 3)).astype(F)).astype(F)   # near or inside the box
    d = np.where((rng.random(n) < 0.75)[:, None], (aim - o).astype(F), rng.uniform(-1, 1, (n, 3)).astype(F)).astype(F)
    z = rng.random((n, 3)) < 0.08
    d[z] = 0
    d[np.all(d == 0, axis=1)] = 1
    onface = rng.random(n) < 0.1
    o[onface, 0] = lo[onface, 0]                         # origin exactly on a slab plane -> 0/0 when d.x == 0
    tmm = np.stack([np.full(n, 0.001, F), np.where(rng.random(n) < 0.5, np.inf, rng.uniform(0.1, 6, n)).astype(F)], 1)
    rays = rayset.as_rays(o, d)
    boxes = np.concatenate([lo, hi], axis=1)
    np.savez_compressed(os.path.join(gold, "aabb_kat.npz"), rays=rays, boxes=boxes, tminmax=tmm,
                        hit=R.aabb_kat(rays, boxes, tmm))


if __name__ == "__main__":
    main()
