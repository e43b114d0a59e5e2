"""N>1 host logic on CPU: two processes (gloo), each "renders" its interleaved row blocks with the host-compiled
device code (tests/devsim -- there is no GPU here), rows are gathered to rank 0 by raytracing-hw_b200/dist.py and the
result must equal the single-process image bit for bit."""
import importlib
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _util


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, S, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "2"
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import _util as u
    rdist = importlib.import_module("raytracing-hw_b200.dist")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    hs = u.rtb.HostScene.load_gltf(u.scene_path("s7_4"), W, H, S)
    p = rdist.rank_params(W, H, S, world, rank, interleave_rows=4)
    rows = rdist.local_rows(p)
    # devsim renders a contiguous row range; render each of this rank's blocks
    local = np.concatenate([u.devsim_render(hs, W, H, S, rows=(int(r), int(r) + 1))["accum"] for r in rows])
    img = rdist.gather_rows(torch.from_numpy(local), W, H, S, world, rank, interleave_rows=4)
    if rank == 0:
        np.save(out_path, img.numpy())
    else:
        assert img is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_equals_single_process(tmp_path):
    W, H, S = 24, 22, 2      # 22 rows: ragged last block, unequal row counts per rank
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(2, _free_port(), W, H, S, out), nprocs=2, join=True)
    hs = _util.rtb.HostScene.load_gltf(_util.scene_path("s7_4"), W, H, S)
    full = _util.devsim_render(hs, W, H, S)["accum"]
    assert np.array_equal(np.load(out), full)


def test_rank_rows_partition():
    rdist = importlib.import_module("raytracing-hw_b200.dist")
    import ctypes as C
    for world in (1, 2, 3, 4, 8):
        seen = []
        for r in range(world):
            p = rdist.rank_params(100, 1080, 4, world, r)
            rows = rdist.local_rows(p)
            n = _util.rtb.lib().rtb_render_local_rows(C.byref(p))
            assert n == len(rows)
            assert [(_util.rtb.lib().rtb_render_global_row(C.byref(p), i)) for i in range(0, n, 37)] == list(rows[::37])
            seen += list(rows)
        assert sorted(seen) == list(range(1080))
