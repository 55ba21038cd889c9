"""Config-5 sweep of the global-memory node-load modes / wavefront tuning in one process:
   python tools/sweep_c5.py <tris> <spheres> <spp> "<tree list: compact,full>" ["slots,threads,th_walk" ...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sexy_raytracer_b200.synth import make_stress_scene
from sexy_raytracer_b200.device import Device
tris, sph, spp = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
lds = sys.argv[4].split(",")       # SRT_WF_TREE values: compact | full
wfs = sys.argv[5:] or ["8388608,1024,16"]
W, H, depth = 1920, 1080, 16
s = make_stress_scene(tris, sph, width=W, height=H, spp=spp, depth=depth)
d = Device(0)
ds = d.upload(s)
P = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=2, seed=7)
ref = None
for wf in wfs:
    os.environ["SRT_WF"] = wf
    for ld in lds:
        os.environ["SRT_WF_TREE"] = ld
        ds.render(P, want_rgba=False)
        out, _, st = ds.render(P, want_rgba=False)
        chk = float(np.nansum(out))
        if ref is None:
            ref = chk
        print("wf=%s tree=%s: %.1f ms %.2f Mseg/s seg=%d sum=%.6e %s" % (wf, ld, st.kernel_ms, st.segments / st.kernel_ms / 1e3, st.segments, chk,
                                                                     "" if abs(chk - ref) <= 1e-4 * abs(ref) else "MISMATCH"), flush=True)
