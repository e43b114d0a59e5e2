"""N>1 host logic on CPU: two processes (gloo), each "renders" its share of the frame -- the rtb_render_params librtb's
rtb_dist_rank_params gives the rank, for both splits -- with the host-compiled device code (tests/devsim: there is no
GPU here); the results are combined on rank 0 by raytracing-hw_b200/dist.py's host mirror of rtb_render_dist's combine
and must equal the single-process frame bit for bit (tile split: the reference's streams; sample split: the same K
streams per pixel, two partial sums added)."""
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


def _worker(rank, world, port, W, H, S, out_path, split=0, K=0):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "2"
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import _util as u
    rdist = importlib.import_module("raytracing-hw_b200.dist")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    hs = u.rtb.HostScene.load_gltf(u.scene_path("s7_4"), W, H, S)
    frame = u.rtb.frame_params(W, H, S, substreams=K, interleave_rows=4)
    p = rdist.rank_params(frame, split, world, rank)
    if split == rdist.SPLIT_TILE:
        local = u.devsim_render(hs, W, H, S, interleave=(p.interleave_count, p.interleave_index, p.interleave_rows), wavefront=True)["accum"]
    else:
        local = u.devsim_render(hs, W, H, S, substreams=p.substreams, substream_range=(p.substream_begin, p.substream_end))["accum"]
    img = rdist.combine_host(torch.from_numpy(local), frame, split, world, rank)
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


def test_two_rank_sample_split_equals_single_process(tmp_path):
    """sample split: rank r renders streams [r K / 2, (r + 1) K / 2) of every pixel; the reduced sums equal the
    single-process K-stream frame when each rank holds ONE stream (a + b has one rounding), and stay within a few ulp
    of it when a rank folds several streams first (the fold order differs: (a + b) + (c + d) vs ((a + b) + c) + d)"""
    W, H, S = 24, 16, 6
    hs = _util.rtb.HostScene.load_gltf(_util.scene_path("s7_4"), W, H, S)
    out = str(tmp_path / "img2.npy")
    mp.spawn(_worker, args=(2, _free_port(), W, H, S, out, 1, 0), nprocs=2, join=True)
    full = _util.devsim_render(hs, W, H, S, substreams=2)["accum"]
    assert np.array_equal(np.load(out), full)
    out4 = str(tmp_path / "img4.npy")
    mp.spawn(_worker, args=(2, _free_port(), W, H, S, out4, 1, 4), nprocs=2, join=True)
    full4 = _util.devsim_render(hs, W, H, S, substreams=4)["accum"]
    assert np.allclose(np.load(out4), full4, rtol=4e-7, atol=0)


def test_rank_params_of_both_splits():
    rtb = _util.rtb
    f = rtb.frame_params(100, 1080, 1024, substreams=0)
    for world in (1, 2, 3, 4, 8):
        streams = []
        for r in range(world):
            q = rtb.rank_params(f, rtb.SPLIT_SAMPLE, world, r)
            assert q.substreams == world and (q.row_begin, q.row_end, q.interleave_count) == (0, 1080, 0)
            streams += list(range(q.substream_begin, q.substream_end))
        assert streams == list(range(world))
    f.substreams = 16
    q = rtb.rank_params(f, rtb.SPLIT_SAMPLE, 8, 3)
    assert (q.substreams, q.substream_begin, q.substream_end) == (16, 6, 8)
    q = rtb.rank_params(f, rtb.SPLIT_TILE, 1, 0)
    assert q.interleave_count == 0 and q.substreams == 16   # K streams per pixel on one GPU: allowed with the tile split


def test_rank_rows_partition():
    rdist = importlib.import_module("raytracing-hw_b200.dist")
    import ctypes as C
    for world in (1, 2, 3, 4, 8):
        seen = []
        for r in range(world):
            p = rdist.rank_params(_util.rtb.frame_params(100, 1080, 4), rdist.SPLIT_TILE, world, r)
            rows = rdist.local_rows(p)
            n = _util.rtb.lib().rtb_render_local_rows(C.byref(p))
            assert n == len(rows)
            assert [(_util.rtb.lib().rtb_render_global_row(C.byref(p), i)) for i in range(0, n, 37)] == list(rows[::37])
            seen += list(rows)
        assert sorted(seen) == list(range(1080))


def test_rank_threads_meet_before_the_collective():
    """rtb_render_multi's rank threads (one process, a thread per GPU) meet before the NCCL step: all leave with the first
    failure of the frame or with success, and a rank that failed before the meeting point does not leave the others waiting
    (csrc/rtb_multisync.h, driven by threads on the host)"""
    import os
    import subprocess
    import _util
    exe = os.path.join(_util.ROOT, "tests", "devsim", "multisync_check")
    subprocess.run(["make", "-C", os.path.dirname(exe), "multisync_check"], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout


def test_block_cache_never_frees_on_a_rank_thread():
    """librtb's block cache (csrc/rtb_block_cache.h) on the host: reuse rule, per-device block limit, byte limit, and the
    deferred frees of rtb_render_multi's rank threads (a cudaFree there can wait for a device whose NCCL kernel waits for
    the freeing rank)"""
    import os
    import subprocess
    import _util
    exe = os.path.join(_util.ROOT, "tests", "devsim", "block_cache_check")
    subprocess.run(["make", "-C", os.path.dirname(exe), "block_cache_check"], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout


def test_multi_gpu_layer_source_on_the_host_with_a_blocking_nccl_stand_in():
    """csrc/rtb_dist.cu itself (rtb_render_multi: a thread per device) compiled for the host, 8 rank threads against a stand-in
    for NCCL that blocks as NCCL does: tile and sample frames complete on rank 0; a rank that fails returns an error instead of
    leaving the others in the collective; no rank thread calls cudaFree however full the block cache is (the 8-GPU hang of r02l);
    and, as the negative control, with the deferral switched off rank threads DO free"""
    import os
    import subprocess
    import _util
    d = os.path.join(_util.ROOT, "tests", "devsim")
    subprocess.run(["make", "-C", d, "dist_emul_check"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    exe, lib = os.path.join(d, "dist_emul_check"), os.path.join(d, "libfakenccl.so")
    r = subprocess.run([exe, lib], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok") and "FAILED" not in r.stdout, r.stdout
    r = subprocess.run([exe, lib], capture_output=True, text=True, timeout=300, env=dict(os.environ, DIST_EMUL_NO_DEFER="1"))
    assert r.returncode != 0 and "FAILED  no cudaFree on a thread that holds a rank" in r.stdout, r.stdout
