#!/usr/bin/env python3
"""Experiment: what would ordering a wavefront's rays by (origin cell, direction octant) buy the closest-hit queue kernel?
Builds bounce-like ray batches on the Sponza stand-in with torch (camera-like rays from the camera position in pixel order,
then one and two diffuse bounces off the hits) and times rtb_closest_hit_device on each batch in pixel order, in a random
permutation and in sorted order at several key granularities.  The answers do not depend on the order (checked by digest).
One JSON line per case."""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import gen_scenes  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")


def morton3(q, bits):
    """interleave the low `bits` bits of the three integer columns of q (torch int64, shape (n, 3))"""
    key = torch.zeros(q.shape[0], dtype=torch.int64, device=q.device)
    for b in range(bits):
        for k in range(3):
            key |= ((q[:, k] >> b) & 1) << (3 * b + k)
    return key


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="sponza")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    path = gen_scenes.ensure(a.workload)
    hs = rtb.HostScene.load_gltf(path, a.width, a.height, 1)
    sc = rtb.Scene(hs, 0)
    cam = hs.camera()
    W, H = a.width, a.height
    n = W * H
    g = torch.Generator(device="cuda")
    g.manual_seed(7)
    # camera-like rays: through the pixel grid (the exact camera arithmetic does not matter here)
    px = (torch.arange(W, device="cuda", dtype=torch.float32) + 0.5) / W * 2 - 1
    py = 1 - (torch.arange(H, device="cuda", dtype=torch.float32) + 0.5) / H * 2
    yy, xx = torch.meshgrid(py, px, indexing="ij")
    axes = torch.from_numpy(cam["axes"]).cuda()
    th = torch.from_numpy(cam["tan_half_fov"]).cuda()
    d = xx[..., None] * th[0] * axes[0] + yy[..., None] * th[1] * axes[1] + axes[2]
    d = (d / d.norm(dim=-1, keepdim=True)).reshape(n, 3)
    o = torch.from_numpy(cam["pos"]).cuda().expand(n, 3)
    rays = torch.cat([o, d], dim=1).contiguous()
    st = torch.cuda.current_stream().cuda_stream
    d_id = torch.empty(n, dtype=torch.int32, device="cuda")
    d_t = torch.empty(n, dtype=torch.float32, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")

    def run(r, name, order=None):
        m = r.shape[0]
        rr = r if order is None else r[order].contiguous()
        d_cnt.zero_()
        sc.closest_hit_device(rr, m, rtb.CULL_GLOBAL, d_id, d_t=d_t, d_counters=d_cnt, stream=st)
        torch.cuda.synchronize()
        box, tri = int(d_cnt[0]), int(d_cnt[1])
        ids = d_id[:m].long()
        if order is not None:
            back = torch.empty_like(ids)
            back[order] = ids
            ids = back
        digest = int((((torch.arange(m, device="cuda") + 1) * (ids + 2)) % 2305843009213693951).sum() % 2305843009213693951)
        for _ in range(2):
            sc.closest_hit_device(rr, m, rtb.CULL_GLOBAL, d_id, d_t=d_t, stream=st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.reps):
            sc.closest_hit_device(rr, m, rtb.CULL_GLOBAL, d_id, d_t=d_t, stream=st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        print(json.dumps({"case": name, "rays": m, "ms": round(ms, 4), "mrays_per_s": round(m / ms * 1e-3, 1), "box_per_ray": round(box / m, 2),
                          "tri_per_ray": round(tri / m, 2), "hit_rate": round(float((ids >= 0).float().mean()), 4), "digest": digest}), flush=True)

    def bounce(r):
        """rays leaving the hits of r: origin just short of the hit, direction uniform on the sphere, flipped to the side the ray came from"""
        m = r.shape[0]
        sc.closest_hit_device(r, m, rtb.CULL_GLOBAL, d_id, d_t=d_t, stream=st)
        torch.cuda.synchronize()
        hit = d_id[:m] >= 0
        t = d_t[:m][hit]
        rr = r[hit]
        p = rr[:, :3] + rr[:, 3:] * (t * 0.9995)[:, None]
        w = torch.randn(p.shape[0], 3, device="cuda", generator=g)
        w = w / w.norm(dim=-1, keepdim=True)
        flip = (w * rr[:, 3:]).sum(-1) > 0
        w = torch.where(flip[:, None], -w, w)
        return torch.cat([p, w], dim=1).contiguous()

    def keys(r, cell_bits, with_octant=True, dir_bits=0):
        lo = r[:, :3].min(0).values
        hi = r[:, :3].max(0).values
        q = ((r[:, :3] - lo) / (hi - lo + 1e-9) * (1 << cell_bits)).long().clamp_(0, (1 << cell_bits) - 1)
        key = morton3(q, cell_bits)
        if with_octant:
            octant = ((r[:, 3] > 0).long() | ((r[:, 4] > 0).long() << 1) | ((r[:, 5] > 0).long() << 2))
            key = (key << 3) | octant
        if dir_bits:
            qd = ((r[:, 3:] * 0.5 + 0.5) * (1 << dir_bits)).long().clamp_(0, (1 << dir_bits) - 1)
            key = (key << (3 * dir_bits)) | morton3(qd, dir_bits)
        return key

    batches = [("camera", rays)]
    b1 = bounce(rays)
    b2 = bounce(b1)
    b3 = bounce(b2)
    batches += [("bounce1", b1), ("bounce2", b2), ("bounce3", b3)]
    for nm, r in batches:
        m = r.shape[0]
        run(r, nm + " pixel order")
        run(r, nm + " shuffled", torch.randperm(m, device="cuda", generator=g))
        if nm == "camera":
            continue
        for cb in (3, 5, 7):
            run(r, nm + " sorted cell %d bits + octant" % cb, torch.argsort(keys(r, cb), stable=True))
        run(r, nm + " sorted octant only (pixel order within)", torch.argsort(keys(r, 0), stable=True))
        run(r, nm + " sorted octant major, cell 5 bits minor", torch.argsort(keys(r, 5, False) | ((keys(r, 0) & 7) << 15), stable=True))
        run(r, nm + " sorted cell 5 bits + octant + dir 2 bits", torch.argsort(keys(r, 5, True, 2), stable=True))
        run(r, nm + " sorted cell 4 bits, dir 3 bits major", torch.argsort(keys(r, 0, False, 3) << 12 | keys(r, 4, False), stable=True))


if __name__ == "__main__":
    main()
