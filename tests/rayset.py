"""Deterministic ray sets for the parity tests (shared by oracle/make_golden.py and tests/)."""
import numpy as np

from sexy_raytracer_b200 import abi

F = np.float32


def as_rays(o, d, time=None):
    r = np.zeros(len(o), abi.RAY_DTYPE)
    r["o"], r["dir"] = np.asarray(o, F), np.asarray(d, F)
    if time is not None:
        r["time"] = np.asarray(time, F)
    return r


def pinhole(scene, step=1):
    """Pixel-centre primary rays, aperture 0, time 0: s=(x+.5)/(W-1), t=((H-y)+.5)/(H-1)."""
    W, H = int(scene.render[0]), int(scene.render[1])
    cam = np.asarray(scene.camera, F)
    origin, lleft, hor, ver = cam[0:3], cam[3:6], cam[6:9], cam[9:12]
    ys, xs = np.mgrid[0:H:step, 0:W:step]
    s = ((xs.ravel().astype(F) + F(0.5)) / F(W - 1)).astype(F)[:, None]
    t = ((F(H) - ys.ravel().astype(F) + F(0.5)) / F(H - 1)).astype(F)[:, None]
    d = (((lleft[None, :] + s * hor[None, :]).astype(F) + (t * ver[None, :]).astype(F)).astype(F) - origin[None, :]).astype(F)
    o = np.broadcast_to(origin, d.shape)
    return as_rays(o, d)


def secondary(hits, n, seed):
    """Random-direction rays leaving reference hit points, random time (exercises the tCur-bounded walk)."""
    rng = np.random.default_rng(seed)
    ok = np.flatnonzero(hits["prim"] >= 0)
    idx = rng.choice(ok, size=n, replace=True)
    d = rng.normal(size=(n, 3)).astype(F)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(F)
    return as_rays(hits["p"][idx], d.astype(F), rng.uniform(0, 1, n).astype(F))


def random_rays(n, seed, lo=-6.0, hi=6.0):
    rng = np.random.default_rng(seed)
    o = rng.uniform(lo, hi, (n, 3)).astype(F)
    o[:, 1] = np.abs(o[:, 1]) + F(0.1)
    tgt = rng.uniform(lo / 2, hi / 2, (n, 3)).astype(F)
    tgt[:, 1] = np.abs(tgt[:, 1])
    return as_rays(o, (tgt - o).astype(F), rng.uniform(0, 1, n).astype(F))


def adversarial(scene, n, seed):
    """Axis-parallel rays (zero direction components -> inf/NaN slabs), rays aimed exactly at triangle
    vertices and edge midpoints, rays starting on a surface, rays along box faces."""
    rng = np.random.default_rng(seed)
    out = []
    k = n // 4
    # axis parallel: one or two zero components
    o = rng.uniform(-4, 4, (k, 3)).astype(F)
    o[:, 1] = np.abs(o[:, 1])
    d = rng.uniform(-1, 1, (k, 3)).astype(F)
    z = rng.integers(0, 3, k)
    d[np.arange(k), z] = 0
    z2 = rng.integers(0, 3, k)
    two = rng.random(k) < 0.5
    d[np.arange(k)[two], z2[two]] = 0
    d[np.all(d == 0, axis=1)] = np.array([0, -1, 0], F)
    out.append(as_rays(o, d, rng.uniform(0, 1, k).astype(F)))
    if scene.n_tri:
        # at vertices / edge midpoints of random triangles, from a fixed eye
        ti = rng.integers(0, scene.n_tri, k)
        eye = np.array([0.5, 3.0, 6.0], F)
        w = rng.integers(0, 3, k)
        v = np.stack([scene.tri_v0[ti], scene.tri_v1[ti], scene.tri_v2[ti]], 1).astype(F)
        a = v[np.arange(k), w]
        b = v[np.arange(k), (w + 1) % 3]
        tgt = np.where((rng.random(k) < 0.5)[:, None], a, ((a + b) * F(0.5)).astype(F))
        out.append(as_rays(np.broadcast_to(eye, tgt.shape), (tgt - eye).astype(F)))
        # starting exactly on a triangle (its centroid), leaving in random directions
        c = ((v[:, 0] + v[:, 1]).astype(F) + v[:, 2]).astype(F) / F(3)
        d = rng.normal(size=(k, 3)).astype(F)
        out.append(as_rays(c.astype(F), d, rng.uniform(0, 1, k).astype(F)))
    # along the faces of BVH boxes: origin below a box corner, direction along a box edge
    nodes = scene.nodes
    ni = rng.integers(0, len(nodes), k)
    o = nodes["bmin"][ni].astype(F) - np.array([0, 0, 3], F)
    d = np.zeros((k, 3), F)
    d[:, 2] = 1
    d[:, 0] = np.where(rng.random(k) < 0.5, 0, rng.uniform(-0.2, 0.2, k)).astype(F)
    out.append(as_rays(o, d))
    return np.concatenate(out)


def aabb_adversarial(n, seed):
    """(rays, boxes [n,6], tminmax [n,2]) aimed at the decision thresholds of the device's box predicates: rays grazing faces,
    edges and corners within a few ulp, exit/entry distances equal to tMax/tMin, |o/d| up to 2^59, zero / denormal / huge
    direction components, degenerate and far-away boxes.  tMin is the walk's constant 0.001 everywhere."""
    rng = np.random.default_rng(seed)
    lo = rng.uniform(-4, 4, (n, 3)).astype(F)
    ext = rng.uniform(0, 3, (n, 3)).astype(F) * (rng.random((n, 3)) > 0.08)
    hi = (lo + ext).astype(F)
    far = rng.random(n) < 0.1                              # boxes a long way from the origin (large |c|)
    shift = (rng.uniform(-1, 1, (n, 3)) * 10.0 ** rng.uniform(2, 6, (n, 1))).astype(F)
    lo[far] = (lo[far] + shift[far]).astype(F); hi[far] = (hi[far] + shift[far]).astype(F)
    hi = np.maximum(lo, hi)
    # a target point on the box surface: each coordinate snapped to lo / hi with high probability -> faces, edges, corners
    u = rng.random((n, 3)).astype(F)
    tgt = (lo + (hi - lo) * u).astype(F)
    snap = rng.random((n, 3))
    tgt = np.where(snap < 0.35, lo, np.where(snap < 0.7, hi, tgt)).astype(F)
    o = (tgt + rng.normal(size=(n, 3)).astype(F) * (10.0 ** rng.uniform(-2, 1.5, (n, 1))).astype(F)).astype(F)
    d = (tgt - o).astype(F)
    # nudge the direction by a few ulp so that the slab values land on either side of the box planes
    k = rng.integers(-3, 4, (n, 3))
    for step in range(1, 4):
        up = k >= step; dn = k <= -step
        d = np.where(up, np.nextafter(d, F(np.inf)), np.where(dn, np.nextafter(d, F(-np.inf)), d)).astype(F)
    kind = rng.random(n)
    # (a) origins far away with tiny direction components: |o/d| up to 2^59
    a = kind < 0.12
    sc = (2.0 ** rng.uniform(10, 30, (n, 1))).astype(F)
    o = np.where(a[:, None], (o * sc).astype(F), o)
    d = np.where(a[:, None], (d / (2.0 ** rng.uniform(0, 29, (n, 1))).astype(F)).astype(F), d)
    # (b) zero, signed-zero, denormal and huge direction components
    b = (kind >= 0.12) & (kind < 0.3)
    comp = rng.integers(0, 3, n)
    val = rng.choice(np.array([0.0, -0.0, 1e-40, -1e-42, 1e-38, 3e38, -2e38, 1e-30, 5e12], F), n)
    d[np.arange(n)[b], comp[b]] = val[b]
    two = b & (rng.random(n) < 0.4)
    d[np.arange(n)[two], (comp[two] + 1) % 3] = rng.choice(np.array([0.0, -0.0, 1e-41], F), int(two.sum()))
    # (c) origin exactly on a box plane (0/0 when that direction component is zero) or inside the box
    c = (kind >= 0.3) & (kind < 0.4)
    o[c, 0] = lo[c, 0]
    ins = (kind >= 0.4) & (kind < 0.45)
    o[ins] = (lo[ins] + (hi[ins] - lo[ins]) * F(0.5)).astype(F)
    d[np.all(d == 0, axis=1)] = np.array([0, 0, 1], F)
    # tMax: inf, random, or exactly the distance of the target point along the ray (t = 1 by construction where d = tgt - o)
    tmax = np.where(rng.random(n) < 0.4, np.inf, rng.uniform(0.01, 6, n)).astype(F)
    one = rng.random(n) < 0.2
    tmax[one] = np.nextafter(F(1), rng.choice(np.array([0, 2], F), int(one.sum()))).astype(F)
    tmm = np.stack([np.full(n, 0.001, F), tmax], 1).astype(F)
    return as_rays(o, d.astype(F), rng.uniform(0, 1, n).astype(F)), np.concatenate([lo, hi], axis=1).astype(F), tmm
