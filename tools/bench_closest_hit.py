#!/usr/bin/env python3
"""cfg3 microbench (BASELINE.json configs[2]): closest hit on a 1 M-triangle soup, random rays, device-resident.
Prints one JSON line per culling mode: Mrays/s, box/triangle tests per ray, algorithmic GB/s (28 B/box + 36 B/tri)."""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import gen_scenes  # noqa: E402
import torch  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--half", type=float, default=10.0)
    ap.add_argument("--rays", type=int, default=16_000_000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--sorted", action="store_true", help="sort rays by origin cell (coherent batch)")
    a = ap.parse_args()
    tri = gen_scenes.soup_triangles(a.tris, seed=1234, half_extent=a.half).reshape(-1, 9)
    hs = rtb.HostScene.from_triangles(tri)
    sc = rtb.Scene(hs, 0)
    rays = gen_scenes.soup_rays(a.rays, seed=42, half_extent=a.half)
    d_rays = torch.from_numpy(rays).cuda()
    d_id = torch.empty(a.rays, dtype=torch.int32, device="cuda")
    d_t = torch.empty(a.rays, dtype=torch.float32, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for cull, nm in ((rtb.CULL_REFERENCE, "reference"), (rtb.CULL_GLOBAL, "global")):
        d_cnt.zero_()
        sc.closest_hit_device(d_rays, a.rays, cull, d_id, d_t=d_t, d_counters=d_cnt, stream=st)
        torch.cuda.synchronize()
        box, trt = int(d_cnt[0]), int(d_cnt[1])
        for _ in range(2):
            sc.closest_hit_device(d_rays, a.rays, cull, d_id, d_t=d_t, stream=st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.reps):
            sc.closest_hit_device(d_rays, a.rays, cull, d_id, d_t=d_t, stream=st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        gbs = (28.0 * box + 36.0 * trt) / (ms * 1e-3) * 1e-9
        print(json.dumps({"workload": "soup %d tris, %d rays" % (a.tris, a.rays), "cull": nm, "ms": ms,
                          "mrays_per_s": a.rays / ms * 1e-3, "box_per_ray": box / a.rays, "tri_per_ray": trt / a.rays,
                          "algorithmic_GBps": gbs, "frac_of_hbm_peak": gbs / peak, "traversal_MB": sc.traversal_bytes / 1e6,
                          "hit_rate": float((d_id >= 0).float().mean()), "host_build_s": hs.build_seconds}))


if __name__ == "__main__":
    main()
