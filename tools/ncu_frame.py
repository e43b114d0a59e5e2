#!/usr/bin/env python3
"""One frame of a bench workload rendered the way bench.py's roofline leg renders it (RTB_RENDER_STAGE_TIMES: one slot group,
every k_wf_trace launch holds all live slots), for `ncu -k regex:k_wf_trace -s N -c 1`: launch N is then a steady-state launch of
the shape `roofline.achieved` is quoted on.  usage: ncu_frame.py --workload sponza --spp 8   (prints the stage times)"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import bench  # noqa: E402  (workload table)
import gen_scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="sponza", choices=sorted(bench.WORKLOADS))
ap.add_argument("--spp", type=int, default=8)
a = ap.parse_args()
import torch  # noqa: E402
rtb = importlib.import_module("raytracing-hw_b200")
scene, W, H, _ = bench.WORKLOADS[a.workload]
hs = rtb.HostScene.load_gltf(gen_scenes.ensure(scene), W, H, a.spp)
sc = rtb.Scene(hs, 0)
p = rtb.frame_params(W, H, a.spp, pixel0_seed=1, flags=rtb.RENDER_STAGE_TIMES)
p = rtb.rank_params(p, rtb.SPLIT_TILE, 1, 0)
d_acc = torch.empty((H, W, 3), dtype=torch.float32, device="cuda")
sc.render_device(p, d_accum=d_acc, d_rgb8=None, d_counters=None, stream=torch.cuda.current_stream().cuda_stream)
trace_ms, other_ms, n = sc.stage_ms()
print(json.dumps({"workload": a.workload, "spp": a.spp, "trace_ms": trace_ms, "other_ms": other_ms, "trace_launches": n,
                  "trace_ms_avg": trace_ms / max(1, n)}))
