"""Multi-GPU driver logic: samples are independent Monte-Carlo estimates, so the
spp range is sharded across ranks with the scene replicated and no data-path collective; the only
exchange is one reduce(sum, float32) of the W*H*3 radiance accumulators onto rank 0 (NCCL over
NVLink on the GPU box; gloo in the CPU tests).  One process per GPU (torch.distributed)."""
from __future__ import annotations


def spp_slice(spp_total: int, rank: int, world: int) -> tuple[int, int]:
    """Strong split: rank k renders GLOBAL sample indices [begin, end) of every pixel; the slices
    tile [0, spp_total) exactly, sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world or spp_total < 0:
        raise ValueError("bad rank/world/spp")
    base, extra = divmod(spp_total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def weak_slice(spp_per_rank: int, rank: int) -> tuple[int, int]:
    """Weak split (bench.py): every rank renders spp_per_rank samples; rank k owns [k*spp, (k+1)*spp)."""
    return rank * spp_per_rank, (rank + 1) * spp_per_rank


def reduce_accumulator(accum, dist=None, dst: int = 0):
    """Sum the per-rank float32 radiance accumulators onto `dst` in place (one collective)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def render_sharded(device_scene, cam, spp_total: int, max_depth: int, accum, dist=None, mode: int = 0, seed: int = 0x3360,
                   pool_size: int = 0):
    """Render this rank's slice of spp_total into `accum` (a CUDA float32 tensor of W*H*3), then reduce
    onto rank 0.  Returns the rank-local stats; rank 0's accum then holds the sum over all spp_total samples."""
    rank = dist.get_rank() if dist is not None and dist.is_initialized() else 0
    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    b, e = spp_slice(spp_total, rank, world)
    stats = device_scene.render_device(cam, accum.data_ptr(), b, e, max_depth, mode=mode, seed=seed, pool_size=pool_size)
    reduce_accumulator(accum, dist)
    return stats


def reduce_adaptive(accum, count, dist=None, dst: int = 0, sumsq=None):
    """Adaptive renders (rtb_render_adaptive_device) exchange one more plane: the per-pixel sample counts
    (SURVEY 8e).  Every rank runs the stop rule on its own slice of the global sample indices; the image is
    sum over ranks of rgb_sum / sum over ranks of count.  `count` travels as int32 (NCCL and gloo both sum it)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
        dist.reduce(count, dst=dst, op=dist.ReduceOp.SUM)
        if sumsq is not None:
            dist.reduce(sumsq, dst=dst, op=dist.ReduceOp.SUM)
    return accum, count


def render_sharded_adaptive(device_scene, cam, max_spp_total: int, max_depth: int, accum, sumsq, count, dist=None, rel_threshold: float = 0.05,
                            min_samples: int = 16, criterion: int = 0, mode: int = 1, seed: int = 0x3360):
    """Rank k takes the global sample indices of its slice of [0, max_spp_total) and stops each pixel by the rule
    evaluated on its own samples (min_samples applies per rank); then sums and counts are reduced onto rank 0.
    accum / sumsq: CUDA float32 W*H*3; count: CUDA int32 W*H (the device writes uint32 values < 2^31)."""
    rank = dist.get_rank() if dist is not None and dist.is_initialized() else 0
    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    b, e = spp_slice(max_spp_total, rank, world)
    stats = device_scene.render_adaptive_device(cam, accum.data_ptr(), sumsq.data_ptr(), count.data_ptr(), b, e, max_depth,
                                                rel_threshold=rel_threshold, min_samples=min_samples, criterion=criterion, mode=mode, seed=seed)
    reduce_adaptive(accum, count, dist, sumsq=sumsq)
    return stats
