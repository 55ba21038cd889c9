"""One config-5 render for ncu: python tools/prof_c5.py <tris> <spheres> <spp> [W H depth]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sexy_raytracer_b200.synth import make_stress_scene
from sexy_raytracer_b200.device import Device
tris, sph, spp = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
W, H, depth = (int(x) for x in sys.argv[4:7]) if len(sys.argv) > 6 else (1920, 1080, 16)
t = time.perf_counter()
s = make_stress_scene(tris, sph, width=W, height=H, spp=spp, depth=depth)
print("built %d nodes in %.1f s" % (len(s.nodes), time.perf_counter() - t), flush=True)
d = Device(0)
ds = d.upload(s)
P = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=2, seed=7)
for _ in range(2):
    _, _, st = ds.render(P, want_rgba=False)
    print("c5 %d tris: %.2f ms %.2f Mseg/s seg=%d launches=%d extend %.1f ms shade %.1f ms" % (
        tris, st.kernel_ms, st.segments / st.kernel_ms / 1e3, st.segments, st.launches, st.extend_ms, st.shade_ms), flush=True)
