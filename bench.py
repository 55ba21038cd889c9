"""bench.py — BASELINE.json's metric: M path-segments/s (and frame time) of the path-tracing loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c4|c3|c2|c1] [--spp S] [--impl reference]

A "step" is one whole frame of the workload (every pixel, all samples).  Default workload = BASELINE config 4:
main.cpp's Master Chief + PBR-globe scene at 1280x720, 5000 spp, depth 4, the samples of every pixel split
across the N ranks (strong scaling; one NCCL sum-reduce of the fp32 radiance buffer per frame, resolve on
rank 0).  One process per GPU; for N > 1 launch under torchrun (see the task contract).

`value`  : whole-job M segments/s with the scene resident in HBM when the timed region starts (device time,
           CUDA events on the launch stream, max over ranks).
`e2e`    : the same metric through the host-buffer C-ABI calls a user makes (srt_scene_create from host arrays +
           render + read-back of the RGBA8 frame), host<->device copies inside the timed region.
`roofline`, `cpu_baseline`, `clocks`, `gpu_launches`: see DESIGN.md "Measurement".

`--impl reference` times the reference's own CPU code for the same path (oracle/_ref, compiled from
/root/reference; else the C port) on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpath-segments/s"
CONFIGS = {  # name -> (scene file, W, H, spp, max_bounce)
    "c1": ("c1_shirley", 400, 225, 16, 10),
    "c2": ("c2_cube", 640, 360, 64, 8),
    "c3": ("c3_masterchief", 426, 240, 1000, 4),
    "c4": ("c3_masterchief", 1280, 720, 5000, 4),
    "c5": (None, 1920, 1080, 256, 16),          # synthetic: built on the fly (sexy_raytracer_b200.synth)
}
WORKLOAD_NAMES = {
    "c1": "Shirley sphere scene 400x225 16spp depth 10 (BASELINE config 1)",
    "c2": "cube.gltf+square.gltf 640x360 64spp depth 8 (BASELINE config 2)",
    "c3": "main.cpp default scene (Master Chief + globe) 426x240 1000spp depth 4 (BASELINE config 3)",
    "c4": "main.cpp default scene (Master Chief + globe) 1280x720 5000spp depth 4, spp split across GPUs (BASELINE config 4)",
    "c5": "synthetic stress: 10M-triangle random mesh + 100k spheres 1920x1080 256spp depth 16 (BASELINE config 5)",
}


def load_workload(args):
    from sexy_raytracer_b200.scene import Scene
    name, W, H, spp, depth = CONFIGS[args.config]
    if args.spp:
        spp = args.spp
    if name is None:
        from sexy_raytracer_b200.synth import make_stress_scene
        t = time.perf_counter()
        s = make_stress_scene(args.tris, args.spheres, width=W, height=H, spp=spp, depth=depth)
        print("[bench] built %s: %d nodes in %.1f s" % (s.name, len(s.nodes), time.perf_counter() - t), file=sys.stderr, flush=True)
        return s, W, H, spp, depth
    s = Scene.load(os.path.join(ROOT, "scenes", name + ".npz"))
    s.render = np.array([W, H, spp, depth], np.int32)
    return s, W, H, spp, depth


# ---- clocks ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ---- CPU baseline: the reference's own code on the host cores ----------------------------------------------------
def _cpu_band(job):
    """One process = one row band (the reference is single threaded and its RNG is a process-wide static)."""
    kind, scene_path, W, H, spp, depth, y0, y1, seed = job
    sys.path.insert(0, ROOT)
    from oracle import refbind
    from sexy_raytracer_b200.scene import Scene
    s = Scene.load(scene_path)
    if kind == "reference":
        R = refbind.RefOracle()
        h = R.from_scene(s)
        t = time.perf_counter()
        _, _, seg = R.render(h, s.cam_args, W, H, spp, depth, s.background, y0=y0, y1=y1, seed=seed, want_rgba=False)
        return seg, time.perf_counter() - t
    P = refbind.PortOracle()
    h = P.from_scene(s)
    t = time.perf_counter()
    _, st = P.render(h, s, s.render_params(width=W, height=H, spp=spp, max_bounce=depth, seed=seed), y0=y0, y1=y1, nthreads=1)
    return int(st.segments), time.perf_counter() - t


def cpu_port_sample(args, scene, seconds_per_core=12.0):
    """Config 5 (one 1.2 GB scene): the C port with one thread per core on a centred band of rows at 1 spp."""
    from oracle import refbind
    name, W, H, spp, depth = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    rows = int(max(4, min(H, round(cores * 0.12e6 * seconds_per_core / 3.0 / W))))
    y0 = (H - rows) // 2
    P = refbind.PortOracle()
    h = P.from_scene(scene)
    t = time.perf_counter()
    _, st = P.render(h, scene, scene.render_params(width=W, height=H, spp=1, max_bounce=depth, seed=1000), y0=y0, y1=y0 + rows,
                     nthreads=cores)
    dt = time.perf_counter() - t
    P.free(h)
    seg = int(st.segments)
    return seg / dt / 1e6, {"kind": "port", "cores": cores, "unit": METRIC,
                            "sample": "%dx%d frame, rows %d..%d, 1 spp (of %d), %d threads of one process; %.2f M segments in %.1f s"
                            % (W, H, y0, y0 + rows, spp, cores, seg / 1e6, dt), "wall_s": round(dt, 2)}


def cpu_frame_sample(args, seconds_per_core=12.0, scene=None):
    """A bounded sample of the workload on all host cores: every core renders its own band of rows of the same
    frame at a reduced spp.  -> (M segments/s, description dict)"""
    if CONFIGS[args.config][0] is None:
        return cpu_port_sample(args, scene, seconds_per_core)
    from oracle import refbind
    kind = "reference" if refbind.RefOracle.available() else "port"
    # load the checker in THIS process before the row-band children are forked: they inherit the mapping, and the driver's
    # record of loaded libraries (taken in the parent) shows which implementation the reference arm ran
    _parent_handle = refbind.RefOracle() if kind == "reference" else refbind.PortOracle()   # noqa: F841
    name, W, H, spp, depth = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    seg_per_spp = W * H * 1.8
    spp_cpu = int(max(1, min(spp, 64, round(0.33e6 * seconds_per_core * cores / seg_per_spp))))
    bands = [(kind, os.path.join(ROOT, "scenes", name + ".npz"), W, H, spp_cpu, depth, H * i // cores, H * (i + 1) // cores, 1000 + i)
             for i in range(cores)]
    bands = [b for b in bands if b[7] > b[6]]
    t = time.perf_counter()
    with mp.get_context("fork").Pool(len(bands)) as pool:
        res = pool.map(_cpu_band, bands)
    wall = time.perf_counter() - t
    seg = sum(r[0] for r in res)
    slowest = max(r[1] for r in res)
    return seg / slowest / 1e6, {"kind": kind, "cores": len(bands), "unit": METRIC,
                                 "sample": "%dx%d frame, all rows, %d spp (of %d), %d row bands = %d processes; %.2f M segments in %.1f s (slowest band)"
                                 % (W, H, spp_cpu, spp, len(bands), len(bands), seg / 1e6, slowest), "wall_s": round(wall, 2),
                                 # the reference as shipped is ONE single-threaded process: the median band's own rate
                                 "one_process": {"value": round(statistics.median(r[0] / r[1] for r in res) / 1e6, 4), "unit": METRIC,
                                                 "note": "median over the bands, each a single-threaded process, all running at once"}}


# ---- the B200 arm ------------------------------------------------------------------------------------------------
def algorithmic_bytes_per_segment(c, segments, n_maps=2):
    """bytes/segment = 32 B + 36 Lt + 40 Ls + 24 A_tri + 3 maps A_tri + 28 (ray), with B, Lt, Ls, A_tri
    per segment MEASURED by the kernel's own counters on this workload (srt_set_counters)."""
    B = c["node_visits"] / segments
    Ls = c["sphere_leaf_tests"] / segments
    Lt = (c["leaf_tests"] - c["sphere_leaf_tests"]) / segments
    A = c["triangle_hits"] / segments
    return 32 * B + 36 * Lt + 40 * Ls + (24 + 3 * n_maps) * A + 28, {"B": B, "Lt": Lt, "Ls": Ls, "A_tri": A}


def flops_per_segment(k):
    """flops/segment = 27 B + 65 Lt + 25 Ls + 175 A_tri + 120 (shade) + 20, un-fused ops"""
    return 27 * k["B"] + 65 * k["Lt"] + 25 * k["Ls"] + 175 * k["A_tri"] + 120 + 20


def run_b200(args):
    import torch
    import torch.distributed as dist
    from sexy_raytracer_b200.device import Device
    from sexy_raytracer_b200.dist import sample_split, reduce_frame

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d needs %d ranks (launch under torch.distributed.run); WORLD_SIZE=%d" % (args.gpus, args.gpus, world))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene, W, H, spp, depth = load_workload(args)
    dev = Device(local)
    stream = torch.cuda.current_stream()
    dev.set_stream(stream.cuda_stream)          # library launches + torch events + NCCL on ONE stream
    ds = dev.upload(scene)
    begin, count = sample_split(spp, world, rank)
    P = scene.render_params(width=W, height=H, spp=spp, max_bounce=depth, sample_begin=begin, sample_count=count, seed=2026,
                            variant=args.variant)
    accum = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")
    rgba = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")    # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        accum.zero_()
        st = ds.render_accumulate(accum.data_ptr(), P) if count > 0 else None
        reduce_frame(accum, dst=0)
        if rank == 0:
            dev.resolve_dev(accum.data_ptr(), W, H, spp, rgba.data_ptr())
        return st

    for _ in range(args.warmup):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    seg = paths = launches = ext_launches = 0
    kern_ms = ext_ms = shd_ms = 0.0
    variant = extend_kernel = 0
    barrier()
    for a, b in ev:
        flush.fill_(1)                           # L2 flush between timed iterations, outside the event pair
        if world > 1:
            dist.barrier()
        a.record(stream)
        st = step()
        b.record(stream)
        if st is not None:
            seg += int(st.segments); paths += int(st.paths); kern_ms += st.kernel_ms
            ext_ms += st.extend_ms; shd_ms += st.shade_ms; ext_launches += st.extend_launches
            launches += st.launches; variant = st.variant; extend_kernel = st.extend_kernel
        launches += 1 if rank == 0 else 0                          # k_resolve
    barrier()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    clk = clocks.stop() if rank == 0 else None
    tot = torch.tensor([float(seg), float(paths)], dtype=torch.float64, device="cuda")
    tmax = torch.tensor([ms, kern_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    seg_all, paths_all = tot.tolist()
    ms_all, kern_ms_max = tmax.tolist()
    value = seg_all / (ms_all * 1e-3) / 1e6
    # the frame of the last timed step (rank 0 holds the reduced sum): CRC32 of the RGBA8 bytes + the total of the linear sums,
    # so that runs at different GPU counts can be compared (Philox is keyed by the global sample index: the N-GPU frame equals
    # the 1-GPU frame up to fp32 summation order - RGBA8 bytes may differ by 1 in a few pixels, the total by ~1e-6 relative)
    frame_crc = sum_total = None
    if rank == 0:
        import zlib
        frame_host = rgba.cpu().numpy()
        frame_crc = "%08x" % (zlib.crc32(frame_host.tobytes()) & 0xffffffff)
        sum_total = float(torch.nan_to_num(accum.double(), nan=0.0, posinf=0.0, neginf=0.0).sum().item())
        if args.save_frame:
            np.save(args.save_frame, frame_host.reshape(H, W, 4))

    # ---- e2e: host buffers through the public C-ABI calls, copies inside the timed region
    h2d = scene.host_bytes()
    d2h = W * H * 4
    dev.set_stream(None)
    e2e_seg = 0
    barrier()
    t0 = time.perf_counter()
    e2e_steps, e2e_kernel_ms, e2e_upload_s, e2e_close_s = [], [], [], []
    for _ in range(args.steps):
        t_step = time.perf_counter()
        ds2 = dev.upload(scene)                                        # H2D: the flattened scene, every step
        e2e_upload_s.append(round(time.perf_counter() - t_step, 4))
        if world == 1:
            _, frame, st = ds2.render(P, want_sum=False, want_rgba=True)   # D2H: the RGBA8 frame
        else:
            dev.set_stream(stream.cuda_stream)
            accum.zero_()
            st = ds2.render_accumulate(accum.data_ptr(), P) if count > 0 else None
            reduce_frame(accum, dst=0)
            if rank == 0:
                dev.resolve_dev(accum.data_ptr(), W, H, spp, rgba.data_ptr())
                frame = rgba.cpu()
            dev.set_stream(None)
        e2e_seg += int(st.segments) if st is not None else 0
        e2e_kernel_ms.append(round(float(st.kernel_ms), 1) if st is not None else None)
        t_close = time.perf_counter()
        ds2.close()                  # srt_scene_destroy INSIDE the timed region: a frame's scene replica is released with it
        e2e_close_s.append(round(time.perf_counter() - t_close, 4))
        e2e_steps.append(round(time.perf_counter() - t_step, 4))
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    se = torch.tensor([float(e2e_seg)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dist.all_reduce(se)
    e2e_value = se.item() / te.item() / 1e6
    # the same call sequence with the scene replica kept on the device between frames (what b200Device does after init; the
    # reference's glDevice::rtFrame re-uploads every frame, which is what `value` above times)
    barrier()
    t0 = time.perf_counter()
    cached_seg = 0
    for _ in range(args.steps):
        if world == 1:
            _, frame, st = ds.render(P, want_sum=False, want_rgba=True)
        else:
            dev.set_stream(stream.cuda_stream)
            accum.zero_()
            st = ds.render_accumulate(accum.data_ptr(), P) if count > 0 else None
            reduce_frame(accum, dst=0)
            if rank == 0:
                dev.resolve_dev(accum.data_ptr(), W, H, spp, rgba.data_ptr())
                frame = rgba.cpu()
            dev.set_stream(None)
        cached_seg += int(st.segments) if st is not None else 0
    barrier()
    tc = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    sc_ = torch.tensor([float(cached_seg)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        dist.all_reduce(sc_)
    e2e_cached = sc_.item() / tc.item() / 1e6

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel.  Per-segment work of the REFERENCE algorithm (B box tests, Lt / Ls
        # triangle / sphere leaf tests) is counted by a short pass of the same frame through an exact-walk kernel (the flat
        # megakernel when the tree is in shared memory: same Philox keys, same paths, node visits = the reference's recursion).
        in_smem = scene.nodes.nbytes + 32 <= 227 * 1024
        dev.set_counters(True)
        stc = ds.render_accumulate(accum.data_ptr(), scene.render_params(width=W, height=H, spp=spp, max_bounce=depth,
                                                                         sample_begin=0, sample_count=max(1, min(8, spp)), seed=2026,
                                                                         variant=1 if in_smem else variant))
        c = dev.last_counters()
        dev.set_counters(False)
        bps_all, k = algorithmic_bytes_per_segment(c, int(stc.segments), n_maps=0 if args.config == "c5" else 2)
        # the traversal kernel's part of SURVEY 8(d)'s figures: node + primitive fetches + the ray / box + leaf tests; the
        # accept-time UV and texel bytes and the shading flops belong to the shade kernel
        bps = 32 * k["B"] + 36 * k["Lt"] + 40 * k["Ls"] + 28
        fl_seg = 27 * k["B"] + 65 * k["Lt"] + 25 * k["Ls"]
        kname = {1: "k_render_mega", 2: "k_wf_extend2" if extend_kernel else "k_wf_extend", 3: "k_render_phased"}.get(variant, "?")
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak_hbm = float(peaks.get("hbm_gbs", 6650.0))
        n_l = max(1, ext_launches)
        seg_rank0 = seg / n_l                                  # rank 0's segments (= rays traced) per launch of the dominant kernel
        launch_s = ext_ms / n_l * 1e-3                         # its average launch duration, CUDA events inside the timed region
        # counters of ONE launch of the dominant kernel in this workload's launch shape from the committed ncu captures
        # (profiles/r2_ncu_summary.json names the report each number was read from; r1_traffic.json for kernels unchanged since)
        ncu = None
        full_size = not args.spp and (args.config != "c5" or (args.tris, args.spheres) == (10_000_000, 100_000))
        for fname in ("r2_ncu_summary.json", "r1_traffic.json"):
            try:
                tr = json.load(open(os.path.join(ROOT, "profiles", fname))).get(args.config)
            except (OSError, ValueError):
                continue
            if tr and tr.get("kernel") == kname and tr.get("node_bytes", extend_kernel) == extend_kernel and full_size:
                ncu = dict(tr, file="profiles/" + fname)
                break
        traffic = None
        if ncu:
            traffic = ncu["dram_bytes_per_launch"]
            if ncu.get("slots_in_that_launch"):               # captured on a full pool: scale to this run's average launch
                traffic = int(traffic * seg_rank0 / ncu["slots_in_that_launch"])
        alg_gbs = bps * seg_rank0 / launch_s / 1e9
        fl = fl_seg * seg_rank0 / launch_s / 1e12
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        fp32_nominal = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        unfused_peak, unfused_src = fp32_nominal / 2, "nominal FFMA peak / 2 (148 SMs x 128 lanes x sm_max_mhz)"
        try:   # the un-fused mul + add ceiling measured on this pool's B200 (tools/fp32_peak.cu)
            mp_ = json.load(open(os.path.join(ROOT, "profiles", "r1_fp32_peak_b200.json")))
            unfused_peak, unfused_src = float(mp_["unfused_mul_add_tflops"]), "measured (tools/fp32_peak.cu, profiles/r1_fp32_peak_b200.json)"
        except (OSError, ValueError, KeyError):
            pass
        common = {"kernel": kname, "traffic": traffic,
                  "per_segment": {a: round(b, 3) for a, b in k.items()}, "flops_per_segment": round(fl_seg, 1),
                  "bytes_per_segment": round(bps, 1), "bytes_per_segment_whole_path": round(bps_all, 1),
                  "segments_per_launch": round(seg_rank0, 1), "launch_ms": round(launch_s * 1e3, 4),
                  "launches_per_step": round(n_l / max(1, args.steps), 1),
                  "share_of_step": {kname: round(ext_ms / max(kern_ms, 1e-9), 4), "k_wf_shade2": round(shd_ms / max(kern_ms, 1e-9), 4)},
                  "algorithmic_gbs": round(alg_gbs, 1),
                  "dram_gbs_measured": round(traffic / launch_s / 1e9, 1) if traffic else None,
                  "frac_streaming": round(traffic / launch_s / 1e9 / peak_hbm, 4) if traffic else None,
                  "traffic_over_algorithmic": round(traffic / (bps * seg_rank0), 4) if traffic else None,
                  "hbm_peak_gbs": peak_hbm, "ncu": {x: ncu[x] for x in ncu if x not in ("dram_bytes_per_launch", "slots_in_that_launch")} if ncu else None}
        if in_smem:
            # tree in shared memory: HBM sees only the ray slots of the wave; the kernel is bound by instruction issue / the FP32 and
            # ALU pipes.  Roof = un-fused fp32 (the reference's arithmetic may not contract to FMA): algorithmic flops of the
            # reference's traversal per second against the measured un-fused mul+add ceiling.
            roofline = dict(common, bound="fp32_unfused", achieved=round(fl, 3), peak=round(unfused_peak, 2), peak_source=unfused_src,
                            unit="TFLOP/s", frac=round(fl / unfused_peak, 4),
                            note="algorithmic flops = 27 B + 65 Lt + 25 Ls per segment (SURVEY 8.4) with B, Lt, Ls of the reference's "
                                 "ordered traversal counted on this frame; k_wf_extend2 reaches the same leaves through a conservative "
                                 "walk (FFMA slab test on outward-rounded boxes), so it executes fewer operations than it is credited with")
        else:
            # tree in HBM / L2: every step of a lane is one unpredictable 32-byte sector; the roof is the random-sector rate of the
            # memory system at this footprint (tools/rand_access_bench.cu), not the streaming peak
            foot = float(scene.nodes.nbytes) / 2 + 64.0 * scene.n_prim       # the compact tree (16 B / node) + leaf-ordered primitives
            peak_sec, src = 43.4, "fallback: 1 GB footprint row of profiles/r1_rand_access_b200.txt"
            try:
                rows = []
                for ln in open(os.path.join(ROOT, "profiles", "r1_rand_access_b200.txt")):
                    f = ln.split()
                    if len(f) >= 7 and f[1] == "MB" and f[2] == "32" and f[3] == "1":
                        rows.append((float(f[0]) * 1048576.0, float(f[5])))
                fit = [r for r in rows if r[0] >= foot] or rows[-1:]
                peak_sec, src = fit[0][1], "measured random 32-byte sector rate at a %d MB footprint (profiles/r1_rand_access_b200.txt)" % (fit[0][0] / 1048576)
            except (OSError, ValueError, IndexError):
                pass
            sectors = (traffic if traffic else bps * seg_rank0) / 32.0 / launch_s / 1e9
            roofline = dict(common, bound="hbm_random_sector", achieved=round(sectors, 2), peak=peak_sec, peak_source=src, unit="G sectors/s",
                            frac=round(sectors / peak_sec, 4), footprint_bytes=int(foot),
                            achieved_source="measured DRAM bytes / 32" if traffic else "ALGORITHMIC bytes / 32 (no ncu capture of this workload shape)",
                            fp32_unfused_frac=round(fl / unfused_peak, 4),
                            note="the walk over the compact 16-byte tree with sorted rays is served partly by the L2 (hit rate ~74 %): DRAM sectors "
                                 "per second are what the HBM delivers, against the measured random-sector ceiling at this footprint; "
                                 "frac_streaming = the same bytes against the streaming peak")
        cpu = None
        if world == 1 and not args.no_cpu:
            v, desc = cpu_frame_sample(args, scene=scene)
            cpu = dict(value=round(v, 4), **desc)
        line = {"metric": METRIC, "value": round(value, 2), "unit": "M segments/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_all / args.steps, 3), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD_NAMES[args.config] + (" [spp overridden: %d]" % spp if args.spp else "")
                           + (" [size overridden: %d triangles, %d spheres]" % (args.tris, args.spheres)
                              if args.config == "c5" and (args.tris, args.spheres) != (10_000_000, 100_000) else ""),
                           "width": W, "height": H, "spp": spp, "max_bounce": depth, "parallelism": "spp-split x%d" % world,
                           "scene": ("synthetic, seeded; tree from the bvh.h-equivalent builder" if args.config == "c5" else
                                     "committed flattening of the reference scene as it loads (globe maps missing upstream -> fallback colour)"),
                           "l2": "256 MiB flush write between timed steps"},
                "frame_time_s": round(ms_all / args.steps / 1e3, 4), "spp_per_s": round(spp / (ms_all / args.steps / 1e3), 1),
                "segments_per_step": seg_all / args.steps, "paths_per_step": paths_all / args.steps,
                "render_kernel_ms_per_step": round(kern_ms_max / args.steps, 3),
                "e2e": {"value": round(e2e_value, 2), "unit": "M segments/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "s_per_step": round(te.item() / args.steps, 4), "s_steps_rank0": e2e_steps,
                        "value_scene_resident": round(e2e_cached, 2),
                        "upload_s_steps_rank0": e2e_upload_s, "kernel_ms_steps_rank0": e2e_kernel_ms, "scene_destroy_s_steps_rank0": e2e_close_s,
                        "path": "srt_scene_create(host SoA) + srt_render(host RGBA8 out) + srt_scene_destroy" if world == 1 else
                                "srt_scene_create(host SoA) + srt_render_accumulate + NCCL reduce + srt_resolve + D2H + srt_scene_destroy"},
                "gpu_launches": launches, "variant": variant, "frame_crc": frame_crc, "sum_total": sum_total, "clocks": clk, "roofline": roofline, "cpu_baseline": cpu}
    ds.close()
    dev.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        emit(line)


# ---- the same frame through the C entry point that owns the whole box: srt_render_group (one process, one host thread per GPU,
# peer copies instead of NCCL).  `python bench.py --backend group --gpus N` - NOT under torch.distributed.run.
def run_group(args):
    import zlib
    if int(os.environ.get("RANK", "0")) != 0:
        return                                   # launched under torchrun by mistake: rank 0 alone drives the box
    from sexy_raytracer_b200.device import DeviceGroup
    scene, W, H, spp, depth = load_workload(args)
    grp = DeviceGroup(list(range(args.gpus)))
    n = len(grp)
    sg = grp.upload(scene)
    P = scene.render_params(width=W, height=H, spp=spp, max_bounce=depth, seed=2026, variant=args.variant)
    for _ in range(args.warmup):
        sg.render(P, want_sum=False, want_rgba=True)
    clocks = ClockSampler(0)
    clocks.start()
    seg = paths = launches = 0
    kern_ms = 0.0
    st = None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        acc, frame, st = sg.render(P, want_sum=False, want_rgba=True)
        seg += int(st.segments); paths += int(st.paths); launches += int(st.launches); kern_ms += st.kernel_ms
    wall = time.perf_counter() - t0
    clk = clocks.stop()
    acc, frame, st = sg.render(P, want_sum=True, want_rgba=True)
    crc = "%08x" % (zlib.crc32(frame.tobytes()) & 0xffffffff)
    sum_total = float(np.nan_to_num(acc.astype(np.float64), nan=0.0, posinf=0.0, neginf=0.0).sum())
    if args.save_frame:
        np.save(args.save_frame, frame)
    # e2e: replicas created, frame rendered, replicas destroyed - every step
    e2e_seg, e2e_steps = 0, []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        t1 = time.perf_counter()
        sg2 = grp.upload(scene)
        _, _, st2 = sg2.render(P, want_sum=False, want_rgba=True)
        sg2.close()
        e2e_seg += int(st2.segments)
        e2e_steps.append(round(time.perf_counter() - t1, 4))
    e2e_s = time.perf_counter() - t0
    sg.close()
    grp.close()
    emit({"metric": METRIC, "value": round(seg / wall / 1e6, 2), "unit": "M segments/s", "n_gpus": n, "steps": args.steps,
          "warmup": args.warmup, "ms_per_step": round(wall / args.steps * 1e3, 3), "higher_is_better": True, "scaling": "strong",
          "vs_baseline": None, "dtype": "f32", "data": "synthetic", "backend": "group",
          "timing": "host clock around srt_render_group (scene replicas resident; the call ends with the RGBA8 frame on the host and "
                    "every GPU synchronised); render_kernel_ms_per_step = CUDA events inside the library, slowest GPU",
          "config": {"workload": WORKLOAD_NAMES[args.config] + (" [spp overridden: %d]" % spp if args.spp else ""),
                     "width": W, "height": H, "spp": spp, "max_bounce": depth, "parallelism": "spp-split x%d, srt_render_group" % n,
                     "l2": "no flush: 8 Mi path slots (0.5 GB) stream through the L2 every wave"},
          "frame_time_s": round(wall / args.steps, 4), "segments_per_step": seg / args.steps, "paths_per_step": paths / args.steps,
          "render_kernel_ms_per_step": round(kern_ms / args.steps, 3),
          "e2e": {"value": round(e2e_seg / e2e_s / 1e6, 2), "unit": "M segments/s", "h2d_bytes_per_step": scene.host_bytes() * n,
                  "d2h_bytes_per_step": W * H * 4, "s_per_step": round(e2e_s / args.steps, 4), "s_steps": e2e_steps,
                  "path": "srt_scene_group_create(host SoA) + srt_render_group(host RGBA8 out) + srt_scene_group_destroy"},
          "gpu_launches": launches, "variant": int(st.variant), "frame_crc": crc, "sum_total": sum_total, "clocks": clk,
          "roofline": None, "cpu_baseline": None})


# ---- the reference arm: the reference's own CPU code, all host cores ----------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene, W, H, spp, depth = load_workload(args)
    vals, desc = [], None
    for i in range(args.warmup + args.steps):
        v, desc = cpu_frame_sample(args, seconds_per_core=5.0, scene=scene)
        if i >= args.warmup:
            vals.append(v)
    value = sum(vals) / len(vals)
    cores = desc["cores"]
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": "M segments/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(desc["wall_s"] * 1e3, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_NAMES[args.config], "width": W, "height": H, "spp": spp, "max_bounce": depth,
                       "parallelism": ("%d host threads of the C port on one band of rows" % cores if desc["kind"] == "port" else
                                       "%d host processes on disjoint row bands (the reference is single threaded)" % cores)},
            "cpu_baseline": dict(value=round(value, 4), **desc),
            "e2e": {"value": round(value, 4), "unit": "M segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# stdout carries exactly ONE line, the JSON result: whatever libraries write to file descriptor 1 meanwhile (NCCL prints its
# version banner there when NCCL_DEBUG is set) is sent to stderr, and the line itself goes to the descriptor saved here.
_RESULT_FD = None


def capture_stdout():
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _RESULT_FD is None:
        sys.stdout.buffer.write(data); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--spp", type=int, default=0, help="override the samples per pixel of the config (not the headline)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--backend", default="ranks", choices=["ranks", "group"],
                    help="ranks: one process per GPU (torch.distributed.run, NCCL reduce); group: ONE process, srt_render_group over --gpus GPUs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--tris", type=int, default=10_000_000, help="config c5: triangles")
    ap.add_argument("--spheres", type=int, default=100_000, help="config c5: spheres")
    ap.add_argument("--variant", type=int, default=0, help="0 auto, 1 flat megakernel, 2 wavefront, 3 phased megakernel")
    ap.add_argument("--save-frame", default="", help="rank 0 writes the RGBA8 frame of the last timed step to this .npy file")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.backend == "group":
        run_group(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
