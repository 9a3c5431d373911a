"""Multi-GPU plumbing: the tree shards by root, one process per GPU.

No collective sits on the data path of `bfs`, the CNN or the logits: each rank expands and
evaluates its own contiguous block of roots.  torch.distributed is used for exactly two things:
gathering the per-root logits / values when the caller wants them on every rank (the only exchange
the path has; A+1 floats per root), and timing (barrier + max over ranks).  Backend-agnostic, so the
same code runs over NCCL on the GPUs and over gloo in the CPU tests.
"""
import torch
import torch.distributed as dist


def shard_range(n_roots, rank, world):
    """Contiguous block of roots owned by `rank`: sizes differ by at most one, order preserved."""
    base, rem = divmod(n_roots, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def gather_rows(local, n_total, group=None):
    """All-gather per-root rows (logits, values) sharded by shard_range -> full (n_total, ...) on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    rank = dist.get_rank(group)
    sizes = [shard_range(n_total, r, world)[1] for r in range(world)]
    pad = max(sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: sizes[rank]] = local
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def max_over_ranks(values, device, group=None):
    """Element-wise max over ranks of a list of floats (device-side timings)."""
    t = torch.tensor(values, dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return [float(v) for v in t]


def whole_job_rate(units_per_rank, world, ms):
    """value = units all ranks processed / the slowest rank's time (weak scaling)."""
    return world * units_per_rank / (ms / 1e3)


def allreduce_gradients(parameters, group=None):
    """Average the gradients of a data-parallel PPO minibatch over ranks (one fp32 sum over
    the ~1.69 M-parameter gradient + beta/alpha per minibatch).  All gradients travel in ONE flat
    bucket (6.7 MB: a single NCCL ring/NVLS launch), parameters without a gradient contribute zeros
    so that every rank reduces the same layout."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    params = [p for p in parameters if p.requires_grad]
    if not params:
        return
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float() for p in params])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat /= dist.get_world_size(group)
    off = 0
    for p in params:
        n = p.numel()
        g = flat[off:off + n].reshape(p.shape).to(p.dtype)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += n


def broadcast_parameters(parameters, src=0, group=None):
    """Make every rank start from rank `src`'s weights (called once before training)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    for p in parameters:
        dist.broadcast(p.data, src=src, group=group)
