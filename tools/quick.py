import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sexy_raytracer_b200.scene import Scene
from sexy_raytracer_b200.device import Device
d = Device(0)
for name, W, H, spp in (("c1_shirley", 400, 225, 16), ("c3_masterchief", 426, 240, 1000), ("c3_masterchief", 1280, 720, 200)):
    s = Scene.load("scenes/%s.npz" % name)
    ds = d.upload(s)
    P = s.render_params(width=W, height=H, spp=spp)
    for it in range(2):
        t = time.time(); acc, rgba, st = ds.render(P); dt = time.time() - t
        print(name, W, H, spp, "kernel_ms %.2f wall %.3f Mseg/s %.1f seg/path %.3f" % (st.kernel_ms, dt, st.segments / st.kernel_ms / 1e3, st.segments / st.paths), flush=True)
    d.set_counters(True)
    P2 = s.render_params(width=W, height=H, spp=max(1, spp // 50))
    acc, rgba, st = ds.render(P2)
    print("  counters: nodes/seg %.2f leaf/seg %.2f exact %s" % (st.node_visits / st.segments, st.leaf_tests / st.segments, d.last_counters()), flush=True)
    d.set_counters(False)
    ds.close()
