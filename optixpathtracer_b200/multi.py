"""Multi-GPU partitioning of a frame: replicas + ONE reduce or gather, no ray exchange.

One process per GPU; `torch.distributed` is plumbing only (NCCL on the GPU box, gloo in the CPU tests).  The scene, BVH and
probe are replicated; what is split is the set of (pixel, sub-frame) pairs, which the reference seeds independently
(`tea<4>(pixel, subframe_index)`, sv4/deviceProgram.cu:407), so every rank reproduces exactly the samples a single GPU would.

  sample split  rank r renders sub-frames r, r+N, ... with cfg.accum_mode = ACCUM_SUM, then one fp32 SUM-reduce of the
                W*H*4 accumulation buffer to rank 0 + resolve (divide, display transform).
  tile split    rank r renders the 32-row bands with (band % N) == r (cfg.tile_rank/tile_count) of the SAME sub-frame;
                the bands are gathered to rank 0.  Bit-identical to the single-GPU frame.
"""
import numpy as np
import torch
import torch.distributed as dist

BAND_ROWS = 32          # must match decodeLaunch() in csrc/shade.cu and oracle/raygen_core.h


def subframes_of_rank(rank, world, n_total):
    """Sub-frame indices rank `rank` renders out of n_total (round-robin, so every prefix is balanced)."""
    return list(range(rank, n_total, world))


def band_rows(rank, world, height):
    """Row indices of the launch rows rank `rank` owns in the tile split."""
    rows = np.arange(height)
    return rows[(rows // BAND_ROWS) % max(world, 1) == rank]


def reduce_sum_to_root(accum, root=0):
    """The sample split's one collective: in-place fp32 SUM of the accumulation buffer onto `root`."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=root, op=dist.ReduceOp.SUM)
    return accum


def gather_bands_to_root(image, root=0):
    """Tile split: `image` is this rank's (H, W[, C]) tensor with only its own bands valid.  Returns the assembled image on
    root (None elsewhere).  Bands are sent as one contiguous block per rank, padded to the largest block."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return image
    rank, world = dist.get_rank(), dist.get_world_size()
    H = image.shape[0]
    counts = [len(band_rows(r, world, H)) for r in range(world)]
    mine = torch.as_tensor(band_rows(rank, world, H), device=image.device, dtype=torch.long)
    block = torch.zeros((max(counts),) + tuple(image.shape[1:]), dtype=image.dtype, device=image.device)
    block[:len(mine)] = image.index_select(0, mine)
    out = [torch.empty_like(block) for _ in range(world)] if rank == root else None
    dist.gather(block, out, dst=root)
    if rank != root:
        return None
    full = image.clone()
    for r in range(world):
        rows = torch.as_tensor(band_rows(r, world, H), device=image.device, dtype=torch.long)
        full.index_copy_(0, rows, out[r][:len(rows)])
    return full
