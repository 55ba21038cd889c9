"""Converged-image golden frames FROM THE REFERENCE (oracle/_ref/libref_oracle.so = /root/reference's own code), at the
sample counts BASELINE's configs name — the second half of north_star's correctness definition:

    c3_masterchief_1k_image.npz : config 3, 426x240 x 1000 spp, depth 4  (main.cpp:176,179-180)
    c4_720p_image.npz           : config 4's geometry, 1280x720 x 256 spp, depth 4  (main.cpp:175-181; the full 5000 spp is
                                  ~3 h of 8 cores per render - the GPU's 5000-spp frame is judged against this one's own A/B
                                  spread instead, tests/test_converged.py)
    c2_cube_full_image.npz      : config 2, 640x360 x 64 spp, depth 8

Each frame is rendered TWICE with independent mt19937 streams (A, B): PSNR(A, B) is the reference's own noise floor.
The reference is single threaded with a process-wide RNG (globals.h:30-35), so every render runs as P processes on
disjoint row bands, band i seeded seed + i; the scene is the reference's own construction
(ref_scene_native -> main.cpp:72-146).

    python oracle/make_golden_converged.py [--only c3|c4|c2] [--procs 8]

Stored (small enough to commit): fp32 per-pixel sums where the frame is small (c3, c2); for 720p the RGB8 frame of A, the
16x16-tile means of the linear sums of A and B, per-channel frame totals, and PSNR(A,B) evaluated here on the full frames.
"""
import argparse
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

F = np.float32
BG = np.array([0.53, 0.81, 0.92], F)     # main.cpp:170
CAM_C3 = [0, 3, 5, 0, 2.5, 0, 0, 1, 0, 70.0, 16.0 / 9.0, 0.1, 10.0, 0, 1.0]
CAM_C2 = [3, 3, 5, 0, 0.5, 0, 0, 1, 0, 40.0, 16.0 / 9.0, 0.0, 10.0, 0, 1.0]
# name -> (native scene id, camera, W, H, spp, depth, output file, keep full sums)
JOBS = {
    "c2": (2, CAM_C2, 640, 360, 64, 8, "c2_cube_full_image.npz", False),
    "c3": (3, CAM_C3, 426, 240, 1000, 4, "c3_masterchief_1k_image.npz", True),
    "c4": (3, CAM_C3, 1280, 720, 256, 4, "c4_720p_image.npz", False),
}
SEED_A, SEED_B = 5489, 20261104


def _band(job):
    which, cam, W, H, spp, depth, y0, y1, seed = job
    from oracle.refbind import RefOracle
    R = RefOracle()
    h = R.native(which)
    acc, rgba, seg = R.render(h, np.array(cam, F), W, H, spp, depth, BG, y0=y0, y1=y1, seed=seed)
    R.free(h)
    return y0, y1, acc[y0:y1].copy(), rgba[y0:y1].copy(), seg


def render(pool, procs, which, cam, W, H, spp, depth, seed):
    # more bands than processes: rows of sky are cheap, rows through the model are not
    nb = procs * 4
    edges = [H * i // nb for i in range(nb + 1)]
    jobs = [(which, cam, W, H, spp, depth, edges[i], edges[i + 1], seed + i) for i in range(nb) if edges[i + 1] > edges[i]]
    acc = np.zeros((H, W, 3), F)
    rgba = np.zeros((H, W, 4), np.uint8)
    seg = 0
    for y0, y1, a, r, s in pool.imap_unordered(_band, jobs):
        acc[y0:y1], rgba[y0:y1] = a, r
        seg += s
    return acc, rgba, seg


def tiles16(x):
    H, W, _ = x.shape
    x = np.where(np.isfinite(x), x, 0)[: H // 16 * 16, : W // 16 * 16].astype(np.float64)
    return x.reshape(H // 16, 16, W // 16, 16, 3).mean(axis=(1, 3)).astype(F)


def psnr_u8(a, b, mask):
    d = (a[..., :3].astype(np.float64) - b[..., :3].astype(np.float64)) ** 2
    mse = d[mask].mean()
    return 99.0 if mse == 0 else float(10 * np.log10(255.0 ** 2 / mse))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--procs", type=int, default=os.cpu_count() or 1)
    args = ap.parse_args()
    gold = os.path.join(ROOT, "tests", "golden")
    with mp.get_context("fork").Pool(args.procs) as pool:
        for name, (which, cam, W, H, spp, depth, fname, full) in JOBS.items():
            if args.only and name != args.only:
                continue
            t0 = time.time()
            a, ra, sa = render(pool, args.procs, which, cam, W, H, spp, depth, SEED_A)
            b, rb, sb = render(pool, args.procs, which, cam, W, H, spp, depth, SEED_B)
            ok = np.isfinite(a).all(axis=2) & np.isfinite(b).all(axis=2)
            out = dict(spp=np.int32(spp), depth=np.int32(depth), segments=np.array([sa, sb], np.int64),
                       psnr_ab=np.float64(psnr_u8(ra, rb, ok)), tiles_a=tiles16(a), tiles_b=tiles16(b),
                       finite=np.packbits(ok), shape=np.array([H, W], np.int32),
                       total_a=np.where(ok[..., None], a, 0).sum(axis=(0, 1), dtype=np.float64),
                       total_b=np.where(ok[..., None], b, 0).sum(axis=(0, 1), dtype=np.float64),
                       seeds=np.array([SEED_A, SEED_B], np.int64))
            if full:
                out.update(sum_a=a, sum_b=b, rgba_a=ra, rgba_b=rb)
            else:
                out.update(rgb_a=ra[..., :3].copy(), rgb_b=rb[..., :3].copy())
            np.savez_compressed(os.path.join(gold, fname), **out)
            print("%s %dx%d x %d spp: %.1f + %.1f M segments, PSNR(A,B) %.2f dB, %.0f s" %
                  (name, W, H, spp, sa / 1e6, sb / 1e6, out["psnr_ab"], time.time() - t0), flush=True)


if __name__ == "__main__":
    main()
