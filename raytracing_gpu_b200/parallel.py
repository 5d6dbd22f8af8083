"""Multi-GPU partitioning: sample-split rendering with ONE collective (a sum-reduce of the float accumulation
buffers).  Paths are independent and the Philox streams are keyed on (pixel, sample, bounce), so the set of samples
is the same for any number of ranks; the only exchange step is the final image sum.

torch.distributed is plumbing here: process group, NCCL reduce over NVLink, device memory for the buffer."""
import numpy as np
import torch
import torch.distributed as dist

TILE_BLOCK_PIXELS = 1024   # RT_TILE_BLOCK_PIXELS (include/rt_b200.h)


def sample_range(spp_total, world, rank):
    """Contiguous sample-index range [first, first+count) of `rank`; the remainder goes to the lowest ranks."""
    base, rem = divmod(int(spp_total), int(world))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def film_order(width, height):
    """Pixel ids (y * width + x) in the renderer's film order: bands of 4 rows, 8x4-pixel tiles inside a band, a narrower
    last tile when width % 8 != 0 (csrc/wavefront.cu film_xy).  Host restatement used to describe tile-split shares."""
    ids = []
    for y0 in range(0, height, 4):
        hb = min(4, height - y0)
        full = width // 8
        for t in range(full):
            ys, xs = np.mgrid[y0:y0 + hb, 8 * t:8 * t + 8]
            ids.append((ys * width + xs).reshape(-1))
        if width % 8:
            ys, xs = np.mgrid[y0:y0 + hb, 8 * full:width]
            ids.append((ys * width + xs).reshape(-1))
    return np.concatenate(ids) if ids else np.zeros(0, np.int64)


def tile_mask(width, height, world, rank):
    """(H, W) bool: pixels rank `rank` owns under tile-split (film blocks b with b % world == rank)."""
    order = film_order(width, height)
    block = np.arange(order.size) // TILE_BLOCK_PIXELS
    mask = np.zeros(width * height, bool)
    mask[order[block % world == rank]] = True
    return mask.reshape(height, width)


def reduce_and_normalise(accum, spp_total, group=None, dst=0):
    """Sum the per-rank accumulation buffers onto `dst` (one reduce) and divide by the total sample count there.
    Returns the mean image on `dst`, None elsewhere.  Float summation order differs with the world size: images
    agree to ~1e-6 relative, not bitwise."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if rank != dst:
        return None
    return accum * (1.0 / float(spp_total))


def render_distributed(render_sums, height, width, spp_total, device, group=None, partition="sample"):
    """render_sums(first, count, accum_tensor, tile_index, tile_count) must ADD the radiance sums of samples
    [first, first+count) of the pixels of tile share (tile_index, tile_count) into accum_tensor (H, W, 3 float32 on
    `device`) — on the GPU that is Scene.render_device(accum.data_ptr(), ..., tile_index=, tile_count=).
    partition "sample": every rank renders all pixels for its sample range; "tile": every rank renders all samples
    for its film blocks.  Either way the buffers are zero where a rank did nothing, so ONE sum-reduce finishes."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    accum = torch.zeros(height, width, 3, dtype=torch.float32, device=device)
    if partition == "tile":
        render_sums(0, spp_total, accum, rank, world)
    elif partition == "sample":
        first, count = sample_range(spp_total, world, rank)
        if count > 0:
            render_sums(first, count, accum, 0, 1)
    else:
        raise ValueError(f"unknown partition {partition!r}")
    return reduce_and_normalise(accum, spp_total, group)
