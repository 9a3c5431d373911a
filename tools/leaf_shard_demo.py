"""One tree, its leaves split over the ranks (BASELINE.json configs[2]: Pong-class, depth 3, one root per step):
every rank expands its leaf block, evaluates it with K4, reduces per first action, all-gathers 2*A floats (NCCL) and
ends with the same SoftTreeMax logits; rank 0 checks them against the unsharded computation and prints the timing.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/leaf_shard_demo.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from softtreemax_b200 import ops, shard
    from softtreemax_b200.engine import Engine
    from softtreemax_b200.roots import make_roots
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    game, depth, n_roots = os.environ.get("GAME", "synth_pong"), int(os.environ.get("DEPTH", 3)), 16
    eng = Engine(game, device=local)
    A = eng.n_actions
    states, stacks = make_roots(eng, n_roots, seed=77)                 # the same roots on every rank
    g = torch.Generator().manual_seed(4)
    shapes = {"cnn.0": (32, 4, 8, 8), "cnn.2": (64, 32, 4, 4), "cnn.4": (64, 64, 3, 3), "linear.0": (512, 3136),
              "action_net": (A, 512), "value_net": (1, 512)}
    params = {}
    for k, shp in shapes.items():
        w = torch.empty(shp)
        torch.nn.init.orthogonal_(w, gain=1.0, generator=g)
        params[k + ".weight"], params[k + ".bias"] = w, torch.zeros(shp[0])
    eng.cnn_set_weights(params)
    val_coef, beta = 0.99 ** depth, 1.0

    def sharded(r):
        blk = shard.expand_leaf_block(eng, states[r], stacks[r], depth, rank, world)
        head = eng.cnn_forward(blk["stacks"])[0] if blk["stacks"].shape[0] else torch.empty((0, A), device=eng.device)
        stat, cnt = shard.partial_action_stats(head, blk["rewards"], blk["first_action"], A, val_coef, beta)
        return shard.combine_action_stats(stat, cnt)

    for r in range(2):
        sharded(r)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    raws = [sharded(r) for r in range(n_roots)]
    e1.record()
    torch.cuda.synchronize()
    ms = shard.max_over_ranks([e0.elapsed_time(e1)], eng.device)[0]
    if rank == 0:
        worst = 0.0
        for r in range(n_roots):                                        # unsharded reference on this rank
            full = eng.bfs(states[r:r + 1], stacks[r:r + 1], depth)
            head = eng.cnn_forward(full["stacks"])[0]
            _, raw, _ = ops.softtreemax_logits_full(eng, head.reshape(1, A ** depth, A), full["rewards"].reshape(1, -1), val_coef, beta,
                                                    1.0, 0.001, False)
            worst = max(worst, float((raw[0] - raws[r]).abs().max()))
        print(json.dumps({"what": "one tree per step, leaves sharded over ranks", "game": game, "A": A, "depth": depth,
                          "leaves_per_tree": A ** depth, "n_gpus": world, "ms_per_tree": ms / n_roots,
                          "trees_per_s": n_roots / (ms / 1e3), "max_abs_diff_vs_unsharded_raw_logits": worst}), flush=True)
        assert worst < 5e-4, worst
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
