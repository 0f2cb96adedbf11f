"""N>1 host logic on CPU: world_size-2 gloo processes exercise the sample partition and the one
sum-reduce per frame (ray_tracing_in_one_weekend_b200/distributed.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ray_tracing_in_one_weekend_b200.distributed import partition_samples, reduce_frame


def test_partition_tiles_the_sample_range():
    for spp in [0, 1, 7, 16, 256, 4097]:
        for world in [1, 2, 3, 4, 8]:
            r = [partition_samples(spp, world, k) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        partition_samples(16, 2, 2)


def _worker(rank, world, port, spp_total, n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # per-sample deterministic "radiance": the frame must not depend on how samples are split
    begin, end = partition_samples(spp_total, world, rank)
    g = torch.Generator().manual_seed(1234)
    per_sample = torch.rand(spp_total, n, generator=g, dtype=torch.float32)
    accum = per_sample[begin:end].sum(0)
    out = reduce_frame(accum, spp_total)
    if rank == 0:
        q.put(out.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduce_matches_single_rank():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    spp_total, n = 37, 3 * 64
    procs = [ctx.Process(target=_worker, args=(r, 2, port, spp_total, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(1234)
    want = torch.rand(spp_total, n, generator=g, dtype=torch.float32).sum(0) / spp_total
    np.testing.assert_allclose(got, want.numpy(), rtol=1e-5)
