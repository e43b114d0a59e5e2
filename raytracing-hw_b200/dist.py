"""Multi-GPU plumbing: one process per GPU (torch.distributed), rows of the canvas dealt out in interleaved blocks.

The reference has no distributed mode (one OpenMP loop over pixels, src/core/scene.cpp:31).  The render shards
trivially by pixel; rows are cut into blocks of `interleave_rows` rows and block b goes to rank b % world, so that
(i) every pixel keeps its own sequential sample stream -- the image does not depend on the number of GPUs -- and
(ii) expensive image regions are spread over all ranks.  There is no data-path collective: the only exchange is the
end-of-frame gather of each rank's resolved rows (RGB8, or the fp32 sums) to rank 0 over NCCL/NVLink (gloo on CPU
in the tests).  Nothing here computes pixels; it only builds rtb_render_params and moves finished rows."""
import numpy as np
import torch
import torch.distributed as dist

from . import RenderParams

DEFAULT_INTERLEAVE_ROWS = 8


def rank_params(width, height, samples, world, rank, pixel0_seed=1, flags=0, interleave_rows=DEFAULT_INTERLEAVE_ROWS):
    """rtb_render_params of `rank`: all rows when world == 1, else the interleaved row blocks of that rank."""
    if world <= 1:
        return RenderParams(width, height, samples, 0, height, pixel0_seed, 0, flags, 0, 0, 0)
    return RenderParams(width, height, samples, 0, height, pixel0_seed, 0, flags, world, rank, interleave_rows)


def local_rows(p):
    """rows rendered by a call with params p, as global row indices (mirror of rtb_render_global_row)."""
    rows = np.arange(p.row_begin, p.row_end)
    if p.interleave_count <= 1:
        return rows
    blk = (rows - p.row_begin) // p.interleave_rows
    return rows[blk % p.interleave_count == p.interleave_index]


def gather_rows(local, width, height, samples, world, rank, group=None, interleave_rows=DEFAULT_INTERLEAVE_ROWS, dst=0):
    """local: (local_rows, width, C) tensor on this rank's device (cuda with NCCL, cpu with gloo).  Returns the
    (height, width, C) image on rank `dst` (None elsewhere).  One all_gather of equally padded blocks."""
    if world <= 1:
        return local
    counts = [len(local_rows(rank_params(width, height, samples, world, r, interleave_rows=interleave_rows))) for r in range(world)]
    pad = max(counts)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    out = torch.empty((world * pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    if rank != dst:
        return None
    image = torch.empty((height,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        rows = torch.as_tensor(local_rows(rank_params(width, height, samples, world, r, interleave_rows=interleave_rows)),
                               device=local.device, dtype=torch.long)
        image[rows] = out[r * pad: r * pad + counts[r]]
    return image
