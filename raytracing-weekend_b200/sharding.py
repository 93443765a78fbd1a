"""Sample sharding over the GPUs of one box (the reference has no multi-GPU code).

One process per GPU.  Rank g of G renders the samples k = g, g+G, g+2G, ... of every pixel into its
own fp32 accumulator (the seed of a sample depends only on (pixel, k), raygen/raygen.cu:129, so the
image does not depend on G up to fp32 summation order); one reduce(sum) to rank 0 per frame combines
them (NCCL over NVLink on GPUs, gloo in the CPU tests of this host logic); rank 0 tonemaps.
"""
import torch
import torch.distributed as dist


def shard_samples(spp, rank, world):
    """(sample_first, sample_stride, sample_count) of `rank`: samples k = rank (mod world), k < spp."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad shard (spp=%d, rank=%d, world=%d)" % (spp, rank, world))
    return rank, world, len(range(rank, spp, world))


def reduce_frame(accum, dst=0):
    """Sum the per-rank accumulators [Ny, Nx, 4] (rgb sums + sample count) onto rank `dst`."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def render_frame_sharded(render_shard, accum, spp, rank=None, world=None, dst=0):
    """One frame: zero `accum`, `render_shard(first, stride, count)` adds this rank's samples into it,
    reduce onto `dst`.  Returns the sample count this rank rendered.  `render_shard` is
    Director.render_accum on a GPU (the C-ABI rtw_render_accum)."""
    if world is None:
        world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
    if rank is None:
        rank = dist.get_rank() if world > 1 else 0
    first, stride, count = shard_samples(spp, rank, world)
    accum.zero_()
    if accum.is_cuda:
        torch.cuda.synchronize(accum.device)
    render_shard(first, stride, count)
    reduce_frame(accum, dst)
    return count
