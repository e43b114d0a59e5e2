"""Rank plumbing for one-process-per-GPU launches (torchrun) around librtb's own multi-GPU layer.

The partition of the frame and the NCCL combine live in the product (csrc/rtb_dist.cu: rtb_comm_*, rtb_render_dist*);
this module only (i) hands rank 0's NCCL unique id to the other ranks through torch.distributed and builds the rank's
rtb.Comm, and (ii) mirrors the combine on HOST tensors over gloo for the CPU tests (tests/test_dist_cpu.py), where no
GPU -- hence no NCCL -- exists.  Nothing here computes pixels.

  tile split   : interleaved row blocks, block b on rank b % world; combine = gather of each rank's compact rows to
                 rank 0, rows put in place (bit-identical to the one-GPU frame)
  sample split : every rank renders the whole canvas with its share of the K engine streams per pixel; combine = sum
                 reduction of the fp32 partial sums to rank 0"""
import numpy as np
import torch
import torch.distributed as dist

from . import SPLIT_SAMPLE, SPLIT_TILE, Comm, frame_params, rank_params, unique_id   # noqa: F401


def make_comm(world, rank, device):
    """Collective: this rank's rtb.Comm (None when world == 1).  torch.distributed must be initialised."""
    if world <= 1:
        return None
    box = [unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return Comm(box[0], world, rank, device)


def local_rows(p):
    """rows rendered by a call with params p, as global row indices (mirror of rtb_render_global_row)."""
    rows = np.arange(p.row_begin, p.row_end)
    if p.interleave_count <= 1:
        return rows
    blk = (rows - p.row_begin) // p.interleave_rows
    return rows[blk % p.interleave_count == p.interleave_index]


def combine_host(local, frame, split, world, rank, group=None, dst=0):
    """The combine of rtb_render_dist on HOST tensors over the current (gloo) process group -- CPU tests only.
    local: this rank's (rows, width, C) result for rank_params(frame, split, world, rank).  Returns the
    (height, width, C) frame on rank `dst`, None elsewhere."""
    if world <= 1:
        return local
    if split == SPLIT_SAMPLE:
        total = local.clone()
        dist.reduce(total, dst=dst, op=dist.ReduceOp.SUM, group=group)
        return total if rank == dst else None
    counts = [len(local_rows(rank_params(frame, split, world, r))) for r in range(world)]
    parts = [torch.empty((c,) + tuple(local.shape[1:]), dtype=local.dtype) for c in counts] if rank == dst else None
    # (gloo's gather wants equal shapes: send / recv the ragged blocks point to point, as the product does with NCCL)
    if rank == dst:
        parts[dst] = local
        reqs = [dist.irecv(parts[r], src=r, group=group) for r in range(world) if r != dst and counts[r] > 0]
        for q in reqs:
            q.wait()
        image = torch.empty((frame.height,) + tuple(local.shape[1:]), dtype=local.dtype)
        for r in range(world):
            image[torch.as_tensor(local_rows(rank_params(frame, split, world, r)), dtype=torch.long)] = parts[r]
        return image
    if counts[rank] > 0:
        dist.send(local.contiguous(), dst=dst, group=group)
    return None
