"""Multi-GPU partitioning: sample-split rendering with ONE collective (a sum-reduce of the float accumulation
buffers).  Paths are independent and the Philox streams are keyed on (pixel, sample, bounce), so the set of samples
is the same for any number of ranks; the only exchange step is the final image sum.

torch.distributed is plumbing here: process group, NCCL reduce over NVLink, device memory for the buffer."""
import torch
import torch.distributed as dist


def sample_range(spp_total, world, rank):
    """Contiguous sample-index range [first, first+count) of `rank`; the remainder goes to the lowest ranks."""
    base, rem = divmod(int(spp_total), int(world))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


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


def render_distributed(render_sums, height, width, spp_total, device, group=None):
    """render_sums(first, count, accum_tensor) must ADD the radiance sums of samples [first, first+count) into
    accum_tensor (H, W, 3 float32 on `device`) — on the GPU that is Scene.render_device(accum.data_ptr(), ...)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    first, count = sample_range(spp_total, world, rank)
    accum = torch.zeros(height, width, 3, dtype=torch.float32, device=device)
    if count > 0:
        render_sums(first, count, accum)
    return reduce_and_normalise(accum, spp_total, group)
