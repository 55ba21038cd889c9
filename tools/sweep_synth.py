"""Synthetic-shape sweep (SURVEY 8.4, config 5 variants): N random spheres / an M-triangle mesh, one B200.
   python tools/sweep_synth.py [spp]   -> one JSON line per shape"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sexy_raytracer_b200.synth import make_stress_scene
from sexy_raytracer_b200.device import Device
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W, H, depth = 1920, 1080, 16
d = Device(0)
shapes = [(0, 1000), (0, 10000), (0, 100000), (10000, 0), (100000, 0), (1000000, 0), (10000000, 0), (10000000, 100000)]
for tris, sph in shapes:
    t = time.perf_counter()
    s = make_stress_scene(tris, sph, width=W, height=H, spp=spp, depth=depth)
    build_s = time.perf_counter() - t
    ds = d.upload(s)
    P = s.render_params(width=W, height=H, spp=spp, max_bounce=depth, seed=7)
    ds.render(P, want_rgba=False)
    _, _, st = ds.render(P, want_rgba=False)
    d.set_counters(True)
    _, _, sc = ds.render(s.render_params(width=W, height=H, spp=1, max_bounce=depth, seed=7, variant=st.variant), want_rgba=False)
    c = d.last_counters()
    d.set_counters(False)
    seg = int(sc.segments)
    B, L, Ls = c["node_visits"] / seg, c["leaf_tests"] / seg, c["sphere_leaf_tests"] / seg
    bps = 32 * B + 36 * (L - Ls) + 40 * Ls + 28
    msegs = st.segments / st.kernel_ms / 1e3
    print(json.dumps({"triangles": tris, "spheres": sph + 16, "nodes": int(len(s.nodes)), "tree_bytes": int(len(s.nodes)) * 32,
                      "build_s": round(build_s, 2), "variant": int(st.variant), "spp": spp, "segments": int(st.segments),
                      "frame_ms": round(st.kernel_ms, 2), "Mseg_per_s": round(msegs, 1), "segments_per_path": round(st.segments / st.paths, 3),
                      "box_tests_per_segment": round(B, 1), "leaf_tests_per_segment": round(L, 2),
                      "algorithmic_bytes_per_segment": round(bps, 0), "algorithmic_GBps": round(bps * msegs / 1e3, 1),
                      "algorithmic_unfused_TFLOPs": round((27 * B + 65 * (L - Ls) + 25 * Ls) * msegs / 1e6, 3)}), flush=True)
    ds.close()
