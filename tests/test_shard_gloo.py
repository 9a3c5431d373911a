"""World-size-2 (and 3) gloo tests of the N>1 host logic, on CPU: root partitioning, the per-root
gather, and the max-over-ranks timing reduction that bench.py uses."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from softtreemax_b200 import shard


def test_shard_range_partitions_exactly():
    for n in (1, 2, 7, 64, 4096, 4097):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                s, c = shard.shard_range(n, r, world)
                seen += list(range(s, s + c))
            assert seen == list(range(n))


def _worker(rank, world, port, n_roots, A):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s, c = shard.shard_range(n_roots, rank, world)
        full = torch.arange(n_roots * A, dtype=torch.float32).reshape(n_roots, A)
        got = shard.gather_rows(full[s:s + c].clone(), n_roots)
        assert torch.equal(got, full), f"rank {rank}: gather order"
        ms = shard.max_over_ranks([10.0 + rank, 5.0 - rank], "cpu")
        assert ms == [10.0 + world - 1, 5.0]
        assert shard.whole_job_rate(100, world, ms[0]) == world * 100 / (ms[0] / 1e3)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_roots", [(2, 9), (2, 4), (3, 10)])
def test_gather_and_timing_over_gloo(world, n_roots):
    port = 29600 + world * 10 + n_roots
    mp.spawn(_worker, args=(world, port, n_roots, 5), nprocs=world, join=True)
