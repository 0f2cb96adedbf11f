"""Multi-GPU host logic: one process per GPU (torch.distributed), scene replicated, samples-per-pixel
partitioned, ONE sum-reduce of the float32 accumulation buffers per frame.

The reference's closest analogue is the mpsc gather of finished pixels (src/main.rs:72,101,114); paths
are independent (src/main.rs:84-106), so no other exchange exists on this path.
PyTorch is plumbing only here: device memory, streams, NCCL (or gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Tuple


def partition_samples(spp_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Global sample indices [begin, end) of every pixel rendered by `rank`; ranges tile [0, spp_total)
    exactly, sizes differ by at most one."""
    if world < 1 or not (0 <= rank < world) or spp_total < 0:
        raise ValueError("bad partition arguments")
    base, extra = divmod(spp_total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def reduce_frame(accum, spp_total: int, dst: int = 0):
    """Sum the per-rank radiance SUMS into rank `dst` and divide by the frame's total spp there
    (`c /= samples_per_pixel`, src/main.rs:98).  `accum` is a flat float32 torch tensor (CUDA for the
    product path, CPU under gloo in tests); modified in place and returned."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
        rank = dist.get_rank()
    else:
        rank = dst
    if rank == dst:
        accum.div_(float(spp_total))
    return accum


def render_frame(scene, width: int, height: int, spp_total: int, accum, *, rank: int = 0, world: int = 1,
                 stream_ptr: int = 0, max_depth: int = 50, seed: int = 95, mode: int = 0):
    """Render this rank's sample range into the CUDA tensor `accum` (zeroed here), reduce, divide on rank 0."""
    begin, end = partition_samples(spp_total, world, rank)
    accum.zero_()
    params = scene.make_params(width, height, spp_total, sample_begin=begin, sample_end=end, spp_total=spp_total,
                               max_depth=max_depth, seed=seed, mode=mode)
    scene.render_device(params, accum.data_ptr(), stream_ptr, want_stats=False)
    return reduce_frame(accum, spp_total)
