#!/usr/bin/env python3
"""N-GPU frame vs 1-GPU frame on real GPUs, one PROCESS per GPU, through librtb's own combine (rtb_render_dist_device
over an rtb_comm = NCCL).  Launch with
     python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_multigpu.py
tile split  : the gathered RGB8 frame must equal a single-GPU render of the whole canvas, every byte (every pixel owns
              its engine stream, so the partition cannot change a pixel);
sample split: the reduced frame must equal the single-GPU render with the same K streams per pixel up to the rounding
              of the final sum (<= 1 level on <= 0.1 % of the bytes)."""
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
import torch.distributed as dist  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")
rdist = importlib.import_module("raytracing-hw_b200.dist")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = rdist.make_comm(world, rank, local)
    cases = [("s7_4", 200, 150, 8), ("sponza_small", 160, 90, 4)]
    if "--full" in sys.argv:
        cases.append(("sponza", 1920, 1080, 8))
    out = {}
    st = torch.cuda.current_stream().cuda_stream
    for scene, W, H, S in cases:
        if rank == 0:
            gen_scenes.ensure(scene)
        dist.barrier()
        path = gen_scenes.ensure(scene)
        hs = rtb.HostScene.load_gltf(path, W, H, S)
        sc = rtb.Scene(hs, local)
        d_rgb = torch.empty((H, W, 3), dtype=torch.uint8, device="cuda") if rank == 0 else None
        res = {"size": [W, H, S]}
        rtb.render_dist_device(sc, comm, rtb.frame_params(W, H, S), rtb.SPLIT_TILE, d_rgb8_root=d_rgb, stream=st)
        torch.cuda.synchronize()
        if rank == 0:
            full = sc.render(W, H, S)["rgb"]
            res["tile_identical_to_single_gpu"] = bool((d_rgb.cpu().numpy() == full).all())
        K = max(world, 2)
        rtb.render_dist_device(sc, comm, rtb.frame_params(W, H, S, substreams=K), rtb.SPLIT_SAMPLE, d_rgb8_root=d_rgb, stream=st)
        torch.cuda.synchronize()
        if rank == 0:
            kfull = sc.render(W, H, S, substreams=K)["rgb"]
            diff = np.abs(d_rgb.cpu().numpy().astype(np.int32) - kfull.astype(np.int32))
            res["sample_max_level_diff"] = int(diff.max())
            res["sample_bytes_differing"] = int(np.count_nonzero(diff))
            res["sample_ok"] = bool(diff.max() <= 1 and np.count_nonzero(diff) <= 1e-3 * diff.size)
            out[scene] = res
            assert res["tile_identical_to_single_gpu"] and res["sample_ok"], (scene, res)
        dist.barrier()
        sc.close()
    if rank == 0:
        print(json.dumps({"world": world, "check": out}))
    if comm is not None:
        comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
