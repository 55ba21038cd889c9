"""A/B timing of the wavefront's launch shapes on one GPU: python tools/wf_ab.py <spp> <x2>:<threads>:<th_walk>[:<fused>[:<shade threads>[:<shade kernel 1|2|3>[:<park 144|160>]]]] [...]
(x2 = 0 / 16 / 32 selects k_wf_extend / k_wf_extend2 with 16- / 32-byte nodes; fused = 1: k_wf_path, one launch per frame).
Config-4 geometry (1280x720, depth 4)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SRT_DEBUG_TUNING"] = "1"
import numpy as np
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device
spp = int(sys.argv[1])
name = os.environ.get("AB_SCENE", "c3_masterchief")
depth = int(os.environ.get("AB_DEPTH", "4"))
d = Device(0)
s = Scene.load("scenes/%s.npz" % name)
ds = d.upload(s)
P = s.render_params(width=1280, height=720, spp=spp, max_bounce=depth, variant=2, seed=7)
for cfg in sys.argv[2:]:
    f = cfg.split(":")
    x2, th, walk = f[:3]
    os.environ["SRT_WF_FUSED"] = f[3] if len(f) > 3 else "0"
    os.environ["SRT_WF_SHADE_THREADS"] = f[4] if len(f) > 4 else "256"
    os.environ["SRT_WF_SHADE"] = f[5] if len(f) > 5 else "2"
    if len(f) > 6: os.environ["SRT_WF_PARK"] = f[6]
    else: os.environ.pop("SRT_WF_PARK", None)
    os.environ["SRT_WF_X2"] = x2
    os.environ["SRT_WF"] = "%s,%s,%s" % (os.environ.get("AB_SLOTS", "8388608"), th, walk)
    ds.render(P, want_sum=False, want_rgba=False)
    best = None
    for _ in range(3):
        out, _, st = ds.render(P, want_sum=True, want_rgba=False)
        if best is None or st.kernel_ms < best.kernel_ms:
            best = st
    st = best
    print("%-14s x2=%-2d fused=%d frame %.2f ms  %.1f Mseg/s  extend %.3f ms/launch (%d launches, %.1f %%)  shade %.3f ms/launch  sum %.6e" %
          (cfg, st.extend_kernel, st.fused, st.kernel_ms, st.segments / st.kernel_ms / 1e3, st.extend_ms / st.extend_launches, st.extend_launches,
           100 * st.extend_ms / st.kernel_ms, st.shade_ms / st.extend_launches, np.nansum(out)), flush=True)
ds.close()
