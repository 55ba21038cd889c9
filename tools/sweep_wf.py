"""GPU tuning sweep for the wavefront variant (2) against the megakernel (1)."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device

cfgs = [("c3_masterchief", 1280, 720, 100, 4), ("c1_shirley", 1280, 720, 32, 10)]
grid = [(1 << 21, 1024, 20)]
for th in (1024, 768, 512):
    for walk in (8, 14, 20, 26):
        grid.append((1 << 21, th, walk))
for slots in (1 << 19, 1 << 20, 1 << 22, 1 << 23):
    grid.append((slots, 1024, 20))
d = Device(0)
for name, W, H, spp, depth in cfgs:
    s = Scene.load("scenes/%s.npz" % name)
    ds = d.upload(s)
    P1 = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=1, seed=7)
    ds.render(P1, want_rgba=False)
    ref, _, st1 = ds.render(P1, want_rgba=False)
    print("%s v1: %.2f ms %.1f Mseg/s" % (name, st1.kernel_ms, st1.segments / st1.kernel_ms / 1e3), flush=True)
    P2 = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=2, seed=7)
    best = None
    seen = set()
    for g in grid:
        if g in seen:
            continue
        seen.add(g)
        os.environ["SRT_WF"] = "%d,%d,%d" % g
        got, _, st = ds.render(P2, want_rgba=False)
        fin = np.isfinite(ref).all(axis=2)
        err = np.abs(got[fin] - ref[fin]).max() / max(1.0, np.abs(ref[fin]).max())
        v = st.segments / st.kernel_ms / 1e3
        print("  wf %-20s %.2f ms %.1f Mseg/s launches %d seg_same=%s err=%.1e" % (os.environ["SRT_WF"], st.kernel_ms, v, st.launches,
              st.segments == st1.segments, err), flush=True)
        if best is None or v > best[0]:
            best = (v, g)
    print("%s best: %.1f Mseg/s %s (v1 %.1f)" % (name, best[0], best[1], st1.segments / st1.kernel_ms / 1e3), flush=True)
    ds.close()
