"""Quick wavefront timing on C4 geometry: python tools/wf_quick.py [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 200
d = Device(0)
for name, W, H, s_, depth in (("c3_masterchief", 1280, 720, spp, 4), ("c1_shirley", 1280, 720, spp // 4, 10)):
    s = Scene.load("scenes/%s.npz" % name)
    ds = d.upload(s)
    for variant in (1, 2):
        P = s.render_params(width=W, height=H, spp=s_, max_bounce=depth, variant=variant, seed=7)
        ds.render(P, want_rgba=False)
        out, _, st = ds.render(P, want_rgba=False)
        print("%s v%d: %.2f ms %.1f Mseg/s launches %d sum %.6e" % (name, variant, st.kernel_ms, st.segments / st.kernel_ms / 1e3, st.launches, np.nansum(out)), flush=True)
    ds.close()
