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


# --------------------------------------------------------------------------------------------------
# Leaf-block partition of ONE tree over several GPUs (secondary partition; BASELINE config 3:
# Pong depth 3, one root).  Every rank expands the first `split` levels (A^split nodes, cheap), keeps its
# contiguous block of those nodes and resumes the expansion from them (Engine.bfs_resume); a leaf's first
# action is known from its ancestor's index.  The only exchange is the all-gather of 2*A floats per rank:
# the per-first-action partial reductions of the leaf logits.
# --------------------------------------------------------------------------------------------------
def split_level(n_actions, depth, world):
    """Smallest level with at least `world` nodes (so that every rank owns at least one), capped at depth."""
    k = 0
    while k < depth and n_actions ** k < world:
        k += 1
    return k


def expand_leaf_block(eng, root_state, root_stack, depth, rank, world, gamma=0.99, frameskip=4, episodic_life=True,
                      clip_reward=True):
    """This rank's block of the leaves of one root's depth-`depth` tree.
    -> dict(stacks, rewards, done, lives, first_action (L_local,) int64, leaf_lo, leaf_hi)."""
    A = eng.n_actions
    k = split_level(A, depth, world)
    lo, cnt = shard_range(A ** k, rank, world)
    per_node = A ** (depth - k)
    if cnt == 0:
        e = torch.empty
        d = eng.device
        return dict(stacks=e((0, 4, 84, 84), dtype=torch.uint8, device=d), rewards=e((0,), dtype=torch.float32, device=d),
                    done=e((0,), dtype=torch.uint8, device=d), lives=e((0,), dtype=torch.int32, device=d),
                    first_action=e((0,), dtype=torch.int64, device=d), leaf_lo=lo * per_node, leaf_hi=lo * per_node)
    rs, rk = root_state.reshape(1, -1), root_stack.reshape(1, 4, 84, 84)
    if k == 0:
        out = eng.bfs(rs, rk, depth, gamma=gamma, frameskip=frameskip, episodic_life=episodic_life, clip_reward=clip_reward)
    else:
        top = eng.bfs(rs, rk, k, gamma=gamma, frameskip=frameskip, episodic_life=episodic_life, clip_reward=False,
                      want_states=True)
        sl = slice(lo, lo + cnt)
        if k == depth:
            rew = top["rewards"][sl]
            out = dict(stacks=top["stacks"][sl], rewards=torch.sign(rew) if clip_reward else rew, done=top["done"][sl],
                       lives=top["lives"][sl])
        else:
            out = eng.bfs_resume(top["states"][sl].contiguous(), top["stacks"][sl].contiguous(), top["rewards"][sl].contiguous(),
                                 top["done"][sl].contiguous(), top["lives"][sl].contiguous(), k, depth - k, gamma=gamma,
                                 frameskip=frameskip, episodic_life=episodic_life, clip_reward=clip_reward)
    leaf_lo = lo * per_node
    idx = torch.arange(leaf_lo, leaf_lo + cnt * per_node, device=eng.device)
    out["first_action"] = idx // (A ** (depth - 1)) if depth >= 1 else torch.zeros_like(idx)
    out["leaf_lo"], out["leaf_hi"] = leaf_lo, leaf_lo + cnt * per_node
    return out


def partial_action_stats(head, rewards, first_action, n_actions, val_coef, beta, cumulative=False):
    """Per-first-action partial reduction of this rank's leaves (actor_critic_ts.py:66-86 restricted to a block).
    head (L_local, A) leaf action-head outputs, rewards (L_local,).  E-mode: (logsumexp of beta * x, count);
    C-mode: (sum of beta * x, count), x = val_coef * head + reward, every leaf contributing its A head outputs."""
    A = n_actions
    x = val_coef * head + rewards[:, None]                                   # (L_local, A)
    stat = torch.full((A,), float("-inf") if not cumulative else 0.0, dtype=torch.float32, device=head.device)
    count = torch.zeros((A,), dtype=torch.float32, device=head.device)
    for a in range(A):
        sel = x[first_action == a]
        if sel.numel() == 0:
            continue
        count[a] = sel.numel()
        stat[a] = (beta * sel).sum() if cumulative else torch.logsumexp(beta * sel.reshape(-1), dim=0)
    return stat, count


def combine_action_stats(stat, count, cumulative=False, group=None):
    """All-gather the partials of every rank and finish the reduction -> raw logits (A,), -inf where no leaf exists
    (actor_critic_ts.py:81,84-86).  Works without a process group (one rank)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        world = dist.get_world_size(group)
        both = torch.stack([stat, count])                                    # (2, A): the whole exchange
        parts = [torch.empty_like(both) for _ in range(world)]
        dist.all_gather(parts, both.contiguous(), group=group)
        stats, counts = torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts])
    else:
        stats, counts = stat[None], count[None]
    total = counts.sum(0)
    if cumulative:
        raw = stats.sum(0) / total
    else:
        raw = torch.logsumexp(stats, dim=0) - torch.log(total)
    return torch.where(total > 0, raw, torch.full_like(raw, float("-inf")))
