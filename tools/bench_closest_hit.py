#!/usr/bin/env python3
"""cfg3 microbench (BASELINE.json configs[2]): closest hit on a 1 M-triangle soup, random rays, device-resident.
Prints one JSON line per culling mode: Mrays/s, box/triangle tests per ray, algorithmic GB/s (28 B/box + 36 B/tri).
Under torchrun (N ranks, one per GPU) the ray batch is cut into N contiguous slices, one per rank, with no data-path
collective (the scene is replicated); time = max over ranks, Mrays/s = all rays / that time; the per-rank hit-id checksums
are summed so that the N-GPU answer can be compared with the one-GPU answer."""
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
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tri = gen_scenes.soup_triangles(a.tris, seed=1234, half_extent=a.half).reshape(-1, 9)
    hs = rtb.HostScene.from_triangles(tri)
    sc = rtb.Scene(hs, local)
    total_rays = a.rays
    lo, hi = total_rays * rank // world, total_rays * (rank + 1) // world
    rays = gen_scenes.soup_rays(total_rays, seed=42, half_extent=a.half)[lo:hi]
    a.rays = hi - lo
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
        # rank-independent digest of the answer: sum of (ray index + 1) * (id + 2) mod 2^61
        idx = torch.arange(lo, hi, dtype=torch.int64, device="cuda")
        agg = torch.stack([torch.tensor(ms, dtype=torch.float64, device="cuda"), torch.tensor(float(box), dtype=torch.float64, device="cuda"),
                           torch.tensor(float(trt), dtype=torch.float64, device="cuda"), (d_id >= 0).sum().double()])
        digest = (((idx + 1) * (d_id.long() + 2)) % 2305843009213693951).sum().reshape(1)
        mx = agg[:1].clone()
        if world > 1:
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(agg, op=dist.ReduceOp.SUM)
            dist.all_reduce(digest, op=dist.ReduceOp.SUM)
        if rank == 0:
            ms, box, trt, hits = float(mx[0]), float(agg[1]), float(agg[2]), float(agg[3])
            gbs = (28.0 * box + 36.0 * trt) / (ms * 1e-3) * 1e-9
            print(json.dumps({"workload": "soup %d tris, %d rays" % (a.tris, total_rays), "n_gpus": world, "cull": nm, "ms": ms,
                              "mrays_per_s": total_rays / ms * 1e-3, "box_per_ray": box / total_rays, "tri_per_ray": trt / total_rays,
                              "algorithmic_GBps": gbs, "frac_of_hbm_peak_per_gpu": gbs / peak / world, "traversal_MB": sc.traversal_bytes / 1e6,
                              "hit_rate": hits / total_rays, "hit_id_digest": int(digest[0]) % 2305843009213693951, "host_build_s": hs.build_seconds}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
