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


def _grad_worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(rank)                                  # different initial weights per rank
        net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))
        beta = torch.nn.Parameter(torch.tensor(1.0 + rank))
        unused = torch.nn.Parameter(torch.zeros(4))              # never gets a gradient on rank 1
        params = list(net.parameters()) + [beta, unused]
        shard.broadcast_parameters(params)
        ref = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))
        torch.manual_seed(0)
        ref0 = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3))
        for a, b in zip(net.parameters(), ref0.parameters()):
            assert torch.equal(a, b), "broadcast must give every rank rank 0's weights"
        assert float(beta) == 1.0
        del ref
        # every rank: its own minibatch; the averaged gradient must equal the full-batch mean gradient
        g = torch.Generator().manual_seed(77)
        x_all, y_all = torch.randn(world * 4, 6, generator=g), torch.randn(world * 4, 3, generator=g)
        x, y = x_all[rank * 4:(rank + 1) * 4], y_all[rank * 4:(rank + 1) * 4]
        loss = ((net(x) * beta - y) ** 2).mean() + (unused.sum() if rank == 0 else 0.0)
        loss.backward()
        shard.allreduce_gradients(params)
        full = ((ref0(x_all) * 1.0 - y_all) ** 2).mean()
        want = torch.autograd.grad(full, list(ref0.parameters()))
        for p, w in zip(net.parameters(), want):
            assert torch.allclose(p.grad, w, atol=1e-6), f"rank {rank}: averaged gradient"
        assert torch.allclose(unused.grad, torch.full((4,), 1.0 / world))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_gradient_allreduce_matches_full_batch_over_gloo():
    mp.spawn(_grad_worker, args=(2, 29688), nprocs=2, join=True)


def _leaf_worker(rank, world, port, A, depth, cumulative):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import policy_ref
        g = torch.Generator().manual_seed(5)
        L = A ** depth
        head = 3 * torch.randn(L, A, generator=g)
        rew = torch.randint(-2, 3, (L,), generator=g).float()
        val_coef, beta = 0.99 ** depth, 1.7
        _, want, _ = policy_ref.softtreemax_logits(head[None], rew[None], A, val_coef, beta, 1.0, 0.001, cumulative)
        # this rank's block of the level-k nodes -> its contiguous block of leaves
        k = shard.split_level(A, depth, world)
        lo, cnt = shard.shard_range(A ** k, rank, world)
        per = A ** (depth - k)
        sl = slice(lo * per, (lo + cnt) * per)
        fa = torch.arange(L)[sl] // (A ** (depth - 1))
        stat, count = shard.partial_action_stats(head[sl], rew[sl], fa, A, val_coef, beta, cumulative)
        raw = shard.combine_action_stats(stat, count, cumulative)
        assert torch.allclose(raw, want[0], atol=2e-5), (rank, raw, want[0])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,A,depth,cumulative", [(2, 6, 3, False), (3, 4, 2, False), (2, 4, 3, True), (3, 9, 2, False)])
def test_leaf_block_partial_logits_over_gloo(world, A, depth, cumulative):
    """One tree split over ranks by leaf blocks: the all-gathered per-first-action partials reproduce the
    full-tree raw logits of the restatement (actor_critic_ts.py:66-86)."""
    mp.spawn(_leaf_worker, args=(world, 29700 + world * 7 + A + depth, A, depth, cumulative), nprocs=world, join=True)


def test_split_level():
    assert shard.split_level(6, 3, 1) == 0 and shard.split_level(6, 3, 2) == 1 and shard.split_level(6, 3, 8) == 2
    assert shard.split_level(4, 2, 64) == 2            # more ranks than leaves: capped, some ranks own nothing
