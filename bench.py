#!/usr/bin/env python3
"""bench.py -- the reference's headline workload on B200: the Sponza stand-in rendered by the sm_100a integrator.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference ...                     (the reference's own OpenMP Scene::render on the host cores)

A "step" is one full render of the workload (BASELINE.json configs[3]: examples/sponza at 1920x1080, 1024 spp --
on the stand-in scene of scenes/gen_scenes.py, because the reference's geometry buffer was never shipped, SURVEY.md
0.3), through librtb's own multi-GPU entry point rtb_render_dist_device (csrc/rtb_dist.cu): one rank per GPU, one NCCL
combine per frame on rank 0.  --split tile (the default at every N): interleaved row blocks, every pixel keeps the
reference's own stream (bit-identical frame), ncclSend/ncclRecv of the RGB8 rows: `value` and `e2e` are like for like
across N ("bit_exact": true).  --split sample: every rank renders the whole canvas with its share of N engine streams
per pixel, ncclReduce of the fp32 sums + resolve on rank 0 (statistically equivalent to the reference's stream, gated
in tests/, not bit-identical); with the default --split auto and N > 1 it is timed behind the headline and reported
as the separately labelled "sample_split" object ("bit_exact": false).  Total work is fixed as N grows: "scaling":
"strong".

value  : Mpix*spp/s of the whole job, scene resident in HBM, device-timed (CUDA events, max over ranks).
e2e    : same metric through the C ABI with HOST buffers: rtb_scene_create (host->device copy of the whole scene) +
         render + device->host copy of the image, every step.
Prints ONE JSON line (rank 0)."""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))

METRIC = "Sponza Mpix*spp/s (Mrays/s alongside) vs host-CPU OpenMP render"
UNIT = "Mpix*spp/s"
WORKLOADS = {
    # name: (scene generator name, width, height, spp)
    "sponza": ("sponza", 1920, 1080, 1024),      # BASELINE.json configs[3], the config the metric is quoted on
    "s7_4": ("s7_4", 1280, 720, 256),            # configs[1]
    "s6_1": ("s6_1", 320, 240, 16),              # configs[0]
    "soup_10m": ("soup_10m", 3840, 2160, 64),    # configs[4]: 10 M triangles, traversal set > L2 (HBM-bound)
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sponza", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (diagnostics only; 0 = the workload's)")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", default="960x540x24", help="WxHxS of the bounded CPU sample (about 10-30 s of host work)")
    ap.add_argument("--megakernel", action="store_true", help="time the first integrator (one thread per pixel) instead of the ray queue")
    ap.add_argument("--split", default="auto", choices=["auto", "tile", "sample"],
                    help="multi-GPU partition (auto: the bit-exact tile split as the headline, the sample split timed behind it when N > 1); see the module docstring")
    ap.add_argument("--substreams", type=int, default=0, help="engine streams per pixel (0: one per rank with --split sample, else 1)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (diagnostic sweeps)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = str(gpu_index)
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 8 and f[0] == self.idx:
                self.rows.append(f)

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = []
        for k, name in ((4, "hw_slowdown"), (5, "hw_thermal_slowdown"), (6, "sw_thermal_slowdown"), (7, "sw_power_cap")):
            if any(r[k].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        mx = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def workload(args):
    scene, w, h, s = WORKLOADS[args.workload]
    w, h, s = args.width or w, args.height or h, args.spp or s
    name = "%s stand-in %dx%d %d spp" % (scene, w, h, s)
    return scene, w, h, s, name


# ------------------------------------------------------------------------------------------------ CPU side
def cpu_reference_render(scene_path, w, h, s, threads=0):
    """The reference's own Scene::render (OpenMP, as shipped) from oracle/_ref, timed around render() only.
    Falls back to nothing: when oracle/_ref is absent the baseline is reported as unavailable."""
    from oracle import ref_oracle
    if not ref_oracle.available():
        return None
    rs = ref_oracle.RefScene.from_gltf(scene_path, w, h, s)
    # all host threads, explicitly: torchrun exports OMP_NUM_THREADS=1 to its workers
    _, sec = rs.render(threads=threads or (os.cpu_count() or 1))
    rs.close()
    return sec


def cpu_baseline(scene_path, sample, rays_per_sample):
    w, h, s = (int(x) for x in sample.split("x"))
    cores = os.cpu_count() or 1
    sec = cpu_reference_render(scene_path, w, h, s)
    if sec is None:
        return {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": "oracle/_ref not present"}
    v = w * h * s / sec * 1e-6
    return {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "%dx%d x %d spp of the same scene, Scene::render() only, OMP threads = %d, %.2f s" % (w, h, s, cores, sec),
            "mrays_per_s": v * rays_per_sample if rays_per_sample else None}


def run_reference_arm(args):
    """--impl reference: the unmodified reference's CPU render of the workload, bounded sample per step, rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import gen_scenes
    scene, w, h, s, name = workload(args)
    path = gen_scenes.ensure(scene)
    cw, ch, cs = (int(x) for x in args.cpu_sample.split("x"))
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_reference_render(path, cw // 2, ch // 2, 1)
    secs = []
    for _ in range(args.steps):
        sec = cpu_reference_render(path, cw, ch, cs)
        if sec is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_oracle.so is not present on this box"}))
            return
        secs.append(sec)
    ms = statistics.mean(secs) * 1e3
    v = cw * ch * cs / (ms * 1e-3) * 1e-6
    sample = "%dx%d x %d spp of the %s scene per step, Scene::render() only, OMP threads = %d" % (cw, ch, cs, scene, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (stand-in scene, seeded generator)",
        "config": {"workload": name, "sample_per_step": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------ B200 arm
def run_b200_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gen_scenes
    rtb = importlib.import_module("raytracing-hw_b200")
    rdist = importlib.import_module("raytracing-hw_b200.dist")

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if rtb.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: librtb.so has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scene, W, H, S, name = workload(args)
    if rank == 0:
        path = gen_scenes.ensure(scene)       # deterministic generator; ranks > 0 wait for the files
    barrier()
    path = gen_scenes.ensure(scene)
    hs = rtb.HostScene.load_gltf(path, W, H, S)
    sc = rtb.Scene(hs, local)
    # the headline is the bit-exact tile split at every N (like for like with N = 1); the sample split, which is only
    # statistically equivalent to the reference's stream, is timed behind it as a labelled secondary figure (N > 1, auto)
    split_name = args.split if args.split != "auto" else "tile"
    split = rtb.SPLIT_SAMPLE if split_name == "sample" else rtb.SPLIT_TILE
    flags = rtb.RENDER_MEGAKERNEL if args.megakernel else 0
    frame = rtb.frame_params(W, H, S, pixel0_seed=1, flags=flags, substreams=args.substreams)
    params = rtb.rank_params(frame, split, world, rank)           # what this rank renders (for the roofline leg)
    comm = rdist.make_comm(world, rank, local)                    # rtb_comm over NCCL: the product's own communicator
    d_rgb = torch.empty((H, W, 3), dtype=torch.uint8, device="cuda") if rank == 0 else None
    d_cnt = torch.zeros(6, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream

    def step(ev=None):
        if ev:
            ev[0].record()
        rtb.render_dist_device(sc, comm, frame, split, d_rgb8_root=d_rgb, d_counters=d_cnt, stream=stream)
        if ev:
            ev[1].record()

    for _ in range(args.warmup):
        flush.zero_()
        step()
    barrier()
    d_cnt.zero_()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sc.launches
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    barrier()
    t_wall = time.perf_counter()
    for k in range(args.steps):
        flush.zero_()
        step(evs[k])
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop()
    launches = sc.launches - launches0
    step_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    t = torch.tensor([step_ms], dtype=torch.float64, device="cuda")
    cnt = d_cnt.clone()
    nl = torch.tensor([launches], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        dist.all_reduce(nl, op=dist.ReduceOp.SUM)
    ms_per_step = float(t[0]) / args.steps
    rays_per_step = int(cnt[0]) / args.steps
    lq_per_step = int(cnt[1]) / args.steps
    value = W * H * S / (ms_per_step * 1e-3) * 1e-6

    # ---- secondary figure (N > 1, --split auto): the sample-range split of the same frame, device-timed the same way
    sample_split = None
    if world > 1 and args.split == "auto":
        n_sec = max(1, min(args.steps, 3))
        for _ in range(2):
            flush.zero_()
            rtb.render_dist_device(sc, comm, frame, rtb.SPLIT_SAMPLE, d_rgb8_root=d_rgb, d_counters=d_cnt, stream=stream)
        barrier()
        evs2 = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(n_sec)]
        for k in range(n_sec):
            flush.zero_()
            evs2[k][0].record()
            rtb.render_dist_device(sc, comm, frame, rtb.SPLIT_SAMPLE, d_rgb8_root=d_rgb, d_counters=d_cnt, stream=stream)
            evs2[k][1].record()
        barrier()
        t2 = torch.tensor([sum(e[0].elapsed_time(e[1]) for e in evs2)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        ms2 = float(t2[0]) / n_sec
        sample_split = {"value": W * H * S / (ms2 * 1e-3) * 1e-6, "unit": UNIT, "ms_per_step": ms2, "steps": n_sec, "warmup": 2, "bit_exact": False,
                        "parallelism": "sample-range split: %d engine streams per pixel, one per rank, whole canvas on every rank; ncclReduce(sum) of the fp32 "
                                       "sums + resolve on rank 0; statistically equivalent to the reference's stream (gated in tests/), NOT bit-identical" % max(world, args.substreams)}

    # ---- e2e: host buffers through the C ABI; scene upload + render + combine + image download inside the timed region
    h_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() if rank == 0 else None
    upload_bytes = sc.upload_bytes
    e2e_value = None

    def e2e_step():
        s2 = rtb.Scene(hs, local)                                   # H2D: nodes, triangles, shading records, textures
        rtb.render_dist(s2, comm, frame, split, h_img.numpy() if rank == 0 else None)      # D2H of the frame inside the call
        s2.close()

    if not args.no_e2e:
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 2))
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        te = torch.tensor([(time.perf_counter() - t0) / n_e2e], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_value = W * H * S / float(te[0]) * 1e-6
    K = max(1, params.substreams)
    if split == rtb.SPLIT_SAMPLE:
        par = ("sample-range split: %d engine stream(s) per pixel, %d per rank, whole canvas on each of %d rank(s); ncclReduce(sum) of the fp32 sums "
               "+ resolve on rank 0 (rtb_render_dist)" % (K, max(1, params.substream_end - params.substream_begin) if K > 1 else 1, world))
    else:
        par = "tile split: row blocks of 8 interleaved over %d rank(s), the reference's own stream per pixel%s; ncclSend/ncclRecv of RGB8 rows to rank 0 (rtb_render_dist)" % (
            world, "" if K <= 1 else " x %d streams" % K)
    return dict(rtb=rtb, sc=sc, hs=hs, path=path, rank=rank, world=world, W=W, H=H, S=S, name=name, value=value,
                ms_per_step=ms_per_step, rays_per_step=rays_per_step, lq_per_step=lq_per_step, scene=scene,
                clocks=clocks, e2e_value=e2e_value, upload_bytes=upload_bytes, t_wall=t_wall, params=params, frame=frame,
                torch=torch, dist=dist, launches=int(nl[0]), stream=stream, megakernel=args.megakernel, parallelism=par, comm=comm,
                sample_split=sample_split, bit_exact=(split == rtb.SPLIT_TILE and K <= 1))


def roofline(r):
    """Dominant kernel: k_wf_trace, the closest-hit queue (k_render with --megakernel).  Algorithmic bytes = 28 B per
    ray/box test + 36 B per ray/triangle test performed by the shipped traversal order on the scene BVH,
    counted by the counting instantiation of the same kernels at a reduced spp on this rank's pixels and scaled by the
    sample ratio (tests per sample are spp-independent).  Kernel time: CUDA events recorded on the launching stream
    around every queue launch of one extra frame of this rank's share (RTB_RENDER_STAGE_TIMES); `achieved` = all queue
    launches' bytes / all queue launches' time = bytes per launch / average launch duration.
    Denominator (two regimes): a traversal set that fits the L2 is up against the L2 -- `peak` = the rate
    at which this GPU gathers 64-byte records at random indices from an L2-resident buffer (rtb_measure_l2, measured in
    this run; the streaming rate is reported beside it); one that does not is up against HBM -- `peak` =
    MEASURED_PEAKS.json hbm_gbs.  frac_hbm is always reported."""
    rtb, sc, p, torch = r["rtb"], r["sc"], r["params"], r["torch"]
    K = max(1, p.substreams)
    s_local = (p.substream_end - p.substream_begin if K > 1 and p.substream_end > 0 else K) * r["S"] // K   # samples per pixel of this rank
    import ctypes
    rows = rtb.lib().rtb_render_local_rows(ctypes.byref(p))
    spp_c = max(1, min(s_local, 4))
    mk = p.flags & rtb.RENDER_MEGAKERNEL
    pc = rtb.RenderParams(p.width, p.height, spp_c, p.row_begin, p.row_end, p.pixel0_seed, 0, 1 | mk,
                          p.interleave_count, p.interleave_index, p.interleave_rows)
    st = sc.render(p.width, p.height, spp_c, params=pc, want_accum=False, want_rgb=False)["stats"]
    scale = s_local / spp_c
    d_acc = torch.empty((rows, p.width, 3), dtype=torch.float32, device="cuda")
    if r["megakernel"]:
        box, tri = st["box_tests"] * scale, st["tri_tests"] * scale
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sc.render_device(p, d_accum=d_acc, d_rgb8=None, d_counters=None, stream=r["stream"])
        e1.record()
        torch.cuda.synchronize()
        kernel, kernel_ms, n_launch, other_ms = "k_render", e0.elapsed_time(e1), 1, 0.0
    else:
        box, tri = (st["box_tests"] - st["light_box_tests"]) * scale, (st["tri_tests"] - st["light_tri_tests"]) * scale
        pt = rtb.RenderParams.from_buffer_copy(p)
        pt.flags |= rtb.RENDER_STAGE_TIMES
        sc.render_device(pt, d_accum=d_acc, d_rgb8=None, d_counters=None, stream=r["stream"])
        kernel_ms, other_ms, n_launch = sc.stage_ms()
        kernel = "k_wf_trace"
    bytes_all = 28.0 * box + 36.0 * tri
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm_peak, hbm_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        hbm_peak, hbm_src = 6650.0, "fallback 6.65 TB/s (of fallback)"
    achieved = bytes_all / (kernel_ms * 1e-3) * 1e-9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):   # one ncu capture per scene, of a whole-canvas launch: only quoted for a rank that renders that canvas
        t = json.load(open(tpath)).get(r["scene"])
        if t and list(t.get("canvas", [])) == [r["W"], r["H"]] and rows == r["H"]:
            traffic = t.get("dram_bytes_per_launch")
    n_samples = rows * r["W"] * s_local
    rays = max(1.0, st["closest_hit_rays"] * scale)
    l2_bytes = torch.cuda.get_device_properties(sc.device).L2_cache_size
    trav = sc.traversal_bytes
    resident = trav <= l2_bytes
    l2 = None
    try:   # the figures the traversal is really up against, measured here, on this GPU
        stream_gbs, gather_gbs = rtb.measure_l2(sc.device, 32 << 20)
        l2 = {"stream_gbs": stream_gbs, "gather64_gbs": gather_gbs, "frac_of_stream": achieved / stream_gbs, "frac_of_gather64": achieved / gather_gbs,
              "how": "rtb_measure_l2: 32 MB buffer, coalesced 16-byte loads / 64-byte records at random indices (2 x LDG.256)"}
        if not resident:
            hs_gbs, hg_gbs = rtb.measure_l2(sc.device, 1 << 30)
            l2["hbm_stream_gbs"], l2["hbm_gather64_gbs"], l2["frac_of_hbm_gather64"] = hs_gbs, hg_gbs, achieved / hg_gbs
    except Exception as e:   # a measurement aid: never fails the bench
        l2 = {"error": str(e)}
    if resident and l2 and "gather64_gbs" in l2:
        bound, peak, src = "l2", l2["gather64_gbs"], "measured in this run: 64-byte records gathered at random indices from a 32 MB (L2-resident) buffer, rtb_measure_l2"
        note = ("traversal set %.1f MB <= L2 %.0f MB: the kernel's fetches are served by the L2 (SURVEY.md 8d regime 1), so frac = achieved / the measured "
                "L2 random-gather rate; frac_hbm is the same bytes over the HBM peak (not the binding limit)" % (trav / 1e6, l2_bytes / 1e6))
    else:
        bound, peak, src = "hbm", hbm_peak, hbm_src
        note = "traversal set %.1f MB > L2 %.0f MB: HBM-bound regime (SURVEY.md 8d regime 2), frac = achieved / measured HBM peak" % (trav / 1e6, l2_bytes / 1e6)
    return {"bound": bound, "l2": l2, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_hbm": achieved / hbm_peak,
            "hbm_peak": hbm_peak, "traffic": traffic,
            "kernel": kernel, "kernel_launches": int(n_launch), "kernel_ms_total": kernel_ms, "kernel_ms_avg": kernel_ms / max(1, n_launch),
            "kernel_share_of_step": kernel_ms / max(1e-9, kernel_ms + other_ms), "other_kernels_ms_total": other_ms,
            "algorithmic_bytes_per_launch": bytes_all / max(1, n_launch), "peak_source": src,
            "box_tests_per_ray": box / rays, "tri_tests_per_ray": tri / rays, "kernel_mrays_per_s": rays / (kernel_ms * 1e-3) * 1e-6,
            "bytes_per_sample": bytes_all / max(1, n_samples), "traversal_bytes": int(trav), "l2_bytes": int(l2_bytes), "note": note}


def main():
    args = parse_args()
    # stdout carries exactly ONE line, the JSON: libraries that write to file descriptor 1 themselves (NCCL prints
    # "NCCL version ..." there when the box sets NCCL_DEBUG=VERSION) are sent to stderr; our line goes to the real stdout
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        run_reference_arm(args)
        return
    r = run_b200_arm(args)
    roof = roofline(r)
    if r["world"] > 1:
        r["dist"].barrier()
    if r["rank"] == 0:
        rays_per_sample = r["rays_per_step"] / (r["W"] * r["H"] * r["S"])
        line = {
            "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": r["world"], "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic (stand-in scene from the seeded generator scenes/gen_scenes.py; the reference's geometry buffers were never shipped)",
            "config": {"workload": r["name"], "parallelism": r["parallelism"],
                       "l2": "256 MB buffer written between timed steps (L2 flush)", "ray_depth": 6},
            "mrays_per_s": r["rays_per_step"] / (r["ms_per_step"] * 1e-3) * 1e-6,
            "light_pdf_mqueries_per_s": r["lq_per_step"] / (r["ms_per_step"] * 1e-3) * 1e-6,
            "rays_per_sample": rays_per_sample,
            "clocks": r["clocks"],
            "e2e": {"value": r["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": int(r["upload_bytes"]) * r["world"],
                    "d2h_bytes_per_step": r["W"] * r["H"] * 3},
            "gpu_launches": int(r["launches"]),
            "roofline": roof,
            # the frame `value` / `e2e` render equals the one-GPU frame (= the reference's per-pixel stream) bit for bit?
            "bit_exact": bool(r["bit_exact"]),
            # outside the timed region (render seconds exclude load / BVH / upload, reported separately)
            "scene_build": dict(seconds=round(r["hs"].build_seconds, 4), triangles=int(r["hs"].desc.n_tris), **r["hs"].build_info()),
        }
        if r["sample_split"]:
            line["sample_split"] = r["sample_split"]
        if r["world"] == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(r["path"], args.cpu_sample, rays_per_sample)
        print(json.dumps(line))
    if r["world"] > 1:
        r["dist"].barrier()
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
