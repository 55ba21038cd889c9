"""GPU tuning sweep for k_render_phased (variant 3) against k_render_mega (variant 1): same frame, bit-identical sums expected."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device

cfgs = [("c3_masterchief", 1280, 720, 100, 4), ("c1_shirley", 1280, 720, 32, 10)]
if len(sys.argv) > 1 and sys.argv[1] == "quick":
    cfgs = cfgs[:1]
d = Device(0)
for name, W, H, spp, depth in cfgs:
    s = Scene.load("scenes/%s.npz" % name)
    ds = d.upload(s)
    P1 = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=1, seed=7)
    ds.render(P1, want_rgba=False)
    ref, _, st1 = ds.render(P1, want_rgba=False)
    print("%s v1: %.2f ms %.1f Mseg/s" % (name, st1.kernel_ms, st1.segments / st1.kernel_ms / 1e3), flush=True)
    P3 = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=3, seed=7)
    best = None
    grid = [(512, 16, 12, 20)]
    for th in (256, 512, 768, 1024):
        for a, b, c in itertools.product((8, 16, 24), (8, 16), (12, 20, 26)):
            grid.append((th, a, b, c))
    seen = set()
    for g in grid:
        if g in seen:
            continue
        seen.add(g)
        os.environ["SRT_TUNE"] = "%d,%d,%d,%d" % g
        got, _, st = ds.render(P3, want_rgba=False)
        same = bool((np.nan_to_num(got) == np.nan_to_num(ref)).all()) and st.segments == st1.segments
        v = st.segments / st.kernel_ms / 1e3
        print("  tune %-18s %.2f ms %.1f Mseg/s same=%s" % (os.environ["SRT_TUNE"], st.kernel_ms, v, same), flush=True)
        if best is None or v > best[0]:
            best = (v, g)
    print("%s best: %.1f Mseg/s %s (v1 %.1f)" % (name, best[0], best[1], st1.segments / st1.kernel_ms / 1e3), flush=True)
    ds.close()
