"""Multi-GPU host logic for one-process-per-GPU launches (torchrun): the scene is replicated, the samples of every
pixel are partitioned over the ranks, and ONE sum-reduce of the float32 accumulation buffers per frame combines them
(SURVEY.md 8e).  The partition and the reduce live in the C ABI (rt_partition_samples, rt_comm_*,
rt_render_reduce*: NCCL inside librtgpu.so); torch.distributed is only the LAUNCHER here — it ferries the 128-byte
NCCL id from rank 0 to the others and provides barriers.  `reduce_frame` is the same frame logic over a
torch.distributed collective, kept for the CPU (gloo) tests of the partition/reduce/divide arithmetic.

The reference's closest analogue is the mpsc gather of finished pixels (src/main.rs:72,101,114); paths are
independent (src/main.rs:84-106), so no other exchange exists on this path.
"""
from __future__ import annotations

from typing import Tuple

from . import api


def partition_samples(spp_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Global sample indices [begin, end) of every pixel rendered by `rank` (rt_partition_samples): the ranges
    tile [0, spp_total) exactly, sizes differ by at most one."""
    if world < 1 or not (0 <= rank < world) or spp_total < 0:
        raise ValueError("bad partition arguments")
    return api.partition_samples(0, spp_total, world, rank)


def make_comm(device: int) -> "api.Comm":
    """The RtComm of this process's rank.  Collective over torch.distributed's default group, which must be
    initialised (any backend): rank 0 creates the NCCL id, the others receive it."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(), dist.get_rank()
    box = [api.Comm.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return api.Comm.from_id(box[0], world, rank, device)


def reduce_frame(accum, spp_total: int, dst: int = 0):
    """Sum the per-rank radiance SUMS into rank `dst` and divide by the frame's total spp there
    (`c /= samples_per_pixel`, src/main.rs:98) over a torch.distributed collective.  `accum` is a flat float32
    torch tensor; modified in place and returned.  (The product path is rt_render_reduce*; this mirrors its
    arithmetic for the gloo tests.)"""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
        rank = dist.get_rank()
    else:
        rank = dst
    if rank == dst:
        accum.div_(float(spp_total))
    return accum


def render_frame(scene, comm, width: int, height: int, spp_total: int, accum, *, stream_ptr: int = 0,
                 max_depth: int = 50, seed: int = 95, mode: int = 0):
    """One frame through rt_render_reduce_device: this rank's share of [0, spp_total) into the CUDA tensor `accum`
    (overwritten), one NCCL reduce, division on rank 0 — all on ONE stream: `stream_ptr`, or torch's current
    stream of the tensor's device when 0 (never the scene's private stream, which torch's ops do not order with)."""
    if not stream_ptr:
        import torch
        stream_ptr = torch.cuda.current_stream(accum.device).cuda_stream or 1  # 0 = legacy default stream = cudaStreamLegacy (0x1)
    params = scene.make_params(width, height, spp_total, sample_begin=0, sample_end=spp_total, spp_total=spp_total,
                               max_depth=max_depth, seed=seed, mode=mode)
    comm.render_reduce_device([scene], params, [accum.data_ptr()], [stream_ptr])
    return accum
