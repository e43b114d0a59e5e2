#!/usr/bin/env python3
"""N-GPU image == 1-GPU image, on real GPUs over NCCL.  Launch with
     python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_multigpu.py
Every rank renders its interleaved row blocks (raytracing-hw_b200/dist.py), rank 0 gathers the rows with one
all_gather over NCCL and compares them with a single-GPU render of the whole canvas: the sums must be identical bits
(every pixel owns its engine stream, so the partition cannot change a pixel)."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import gen_scenes  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")
rdist = importlib.import_module("raytracing-hw_b200.dist")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {}
    for scene, W, H, S in (("s7_4", 200, 150, 8), ("sponza_small", 160, 90, 4)):
        if rank == 0:
            gen_scenes.ensure(scene)
        dist.barrier()
        path = gen_scenes.ensure(scene)
        hs = rtb.HostScene.load_gltf(path, W, H, S)
        sc = rtb.Scene(hs, local)
        p = rdist.rank_params(W, H, S, world, rank)
        n_local = len(rdist.local_rows(p))
        d_acc = torch.empty((n_local, W, 3), dtype=torch.float32, device="cuda")
        sc.render_device(p, d_accum=d_acc, d_rgb8=None, d_counters=None, stream=torch.cuda.current_stream().cuda_stream)
        img = rdist.gather_rows(d_acc, W, H, S, world, rank)
        if rank == 0:
            full = sc.render(W, H, S)["accum"]
            same = bool((img.cpu().numpy() == full).all())
            out[scene] = {"size": [W, H, S], "identical_to_single_gpu": same}
            assert same, scene
        dist.barrier()
    if rank == 0:
        print(json.dumps({"world": world, "check": out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
