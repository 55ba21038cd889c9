"""One render call for ncu: python tools/prof_one.py <scene> <W> <H> <spp> <depth> <variant>"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device
name, W, H, spp, depth, variant = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
d = Device(0)
s = Scene.load("scenes/%s.npz" % name)
ds = d.upload(s)
P = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, variant=variant, seed=7)
for _ in range(2):
    _, _, st = ds.render(P, want_rgba=False)
print("%s v%d: %.2f ms %.1f Mseg/s seg=%d" % (name, variant, st.kernel_ms, st.segments / st.kernel_ms / 1e3, st.segments), flush=True)
