#!/usr/bin/env python
"""bench.py — the hot path of SoftTreeMax on B200: breadth-first tree expansion (K1 clone, K2 Atari
step, K3 frames) of R root states to depth d.

Contract: one JSON line on stdout (rank 0).  `value` = whole-job leaf states per second with the
roots already resident in HBM; `e2e` = the same through the host-buffer C-ABI call
(stm_bfs_host: roots from pinned host memory, rewards/done/lives back to the host, leaf stacks left
on the GPU as the reference's bfs() leaves them); `roofline` = the dominant kernel (K2), timed live
with CUDA events on the launch stream; `cpu_baseline` = the CPU oracle on the box's host cores on a
bounded sample.  `--impl reference` times the reference's CPU path: the reference's own emulator
(torchcule) is not installable here, so that arm is the oracle port on all host threads.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--roots R] [--depth D] [--impl reference]
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GAME = "synth_breakout"          # default cartridge; --game overrides
STANDS_FOR = {"synth_breakout": "BreakoutNoFrameskip-v4", "synth_pong": "PongNoFrameskip-v4",
              "synth_pacman": "MsPacmanNoFrameskip-v4", "synth_torture": "opcode/TIA torture cartridge"}
_JSON_OUT = None                  # where the JSON line goes (stdout; a private copy of it when NCCL logs share the descriptor)
METRIC = "tree_leaf_states_per_sec"
UNIT = "leaves/s"
CYCLES_PER_NODE_STEP = 4 * 262 * 76          # emulated CPU cycles per node-step (NTSC nominal)
# From the `ncu --set full` capture of both K2 launches of a depth-2 step of 4 096 synth_breakout roots (16 384 + 65 536 envs),
# summarised in profiles/r02_k2_final_ncu_full_summary.txt:
#   DRAM bytes per env   = (dram__bytes_read.sum + dram__bytes_write.sum) / envs: what reaches HBM of the native rows (distinct rows
#                          only, most of them still in L2 when K3 reads them) + the 512 B state trip;
#   warp instructions per env (node-step) = smsp__inst_executed.sum / envs: what K2's issue-slot roofline is computed from.
K2_NCU_DRAM_BYTES_PER_ENV = (0.004644e9 + 0.041460e9 + 0.019603e9 + 0.321036e9) / 81920
K2_NCU_WARP_INST_PER_ENV = (265868543 + 907441145) / 81920
K2_NCU_ISSUE_ACTIVE = {"16384_envs": 0.1516, "65536_envs": 0.2805}       # smsp__issue_active of the two launches


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        self.nv = self.h = None
        try:                                                        # NVML is brought up here, before the timed region
            import pynvml as nv
            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception as e:                                      # NVML missing: report that, do not fail the run
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def run(self):
        nv, h = self.nv, self.h
        if nv is None:
            return
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        try:
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.01)
        except Exception as e:
            self.reasons.add(f"nvml_error:{type(e).__name__}")

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_oracle_rate(depth, seconds, threads, game=None):
    """leaves/s of the CPU oracle (oracle/, the timed CPU baseline) on a bounded sample."""
    from oracle import orc
    g = orc.Game(game or GAME)
    R = 64
    roots, stacks = orc.make_roots(g, R, seed=1234)
    g.bfs(roots[:8], stacks[:8], depth, nthreads=threads)           # warm-up
    leaves, t0 = 0, time.perf_counter()
    while True:
        g.bfs(roots, stacks, depth, want_state=False, nthreads=threads)
        leaves += R * g.n_actions ** depth
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    return leaves / dt, leaves, dt


def run_reference(args, rank):
    """--impl reference: the reference's CPU path.  torchcule (CuLE's CPU back-end, what
    cule_bfs.py runs below its crossover depth) is unavailable on this box, so this is the oracle
    port (oracle/stm_oracle.c, OpenMP over envs) on all host threads, same workload definition."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    vals = []
    for i in range(args.warmup + args.steps):
        rate, leaves, dt = cpu_oracle_rate(args.depth, seconds=max(1.0, 20.0 / max(1, args.steps)), threads=threads, game=args.game)
        if i >= args.warmup:
            vals.append((rate, leaves, dt))
    rate = sum(v[1] for v in vals) / sum(v[2] for v in vals)
    ms = 1e3 * sum(v[2] for v in vals) / len(vals)
    sample = f"{vals[0][1]} leaves per step (64 roots x depth {args.depth}, repeated ~{vals[0][2]:.1f}s), oracle port, OpenMP"
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": dict(workload_config(args), timed_sample=sample,
                           note="a rate: the CPU arm times 64 roots of this workload repeatedly, not roots_per_gpu roots per step"),
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference CuLE CPU back-end unavailable on this box; oracle port timed instead"}
    print(json.dumps(line), flush=True)


def n_actions_of(game):
    import json as _json
    table = _json.load(open(os.path.join(ROOT, "softtreemax_b200", "roms", "games.json")))
    return len(table[game]["actions"])


def workload_config(args):
    A = n_actions_of(args.game)
    return {"workload": f"{args.game} (stands for {STANDS_FOR.get(args.game, args.game)}, A={A}) tree_depth={args.depth} "
                        f"roots_per_gpu={args.roots} frameskip=4 gamma=0.99 clip_reward episodic_life",
            "roots_per_gpu": args.roots, "depth": args.depth, "leaves_per_gpu": args.roots * A ** args.depth,
            "roots": f"root i of rank r = game seed {args.seed} + ((r + i // 192) mod {args.games}), state after 64 + (i mod 192) agent steps "
                     f"of random play (softtreemax_b200/roots.py): every rank expands the same mix of {args.games} games",
            "l2": "working set > L2: leaf stacks written per step = %.2f GB per GPU" % (args.roots * A ** args.depth * 28224 / 1e9),
            "parallelism": f"roots sharded over {args.gpus} GPU(s), no data-path collective"}


def _sb3_params(A, seed=4):
    """Random-init weights of the architecture: orthogonal, SB3 gains (reference default seed 4, utils.py:67)."""
    import math
    import torch
    g = torch.Generator().manual_seed(seed)
    shapes = {"cnn.0": (32, 4, 8, 8), "cnn.2": (64, 32, 4, 4), "cnn.4": (64, 64, 3, 3), "linear.0": (512, 3136),
              "action_net": (A, 512), "value_net": (1, 512)}
    params = {}
    for k, shp in shapes.items():
        w = torch.empty(shp)
        torch.nn.init.orthogonal_(w, gain={"action_net": 0.01, "value_net": 1.0}.get(k, math.sqrt(2)), generator=g)
        params[k + ".weight"], params[k + ".bias"] = w, torch.zeros(shp[0])
    return params


def run_leafshard(args, eng, rank, world, local, steps):
    """BASELINE.json configs[2]: PongNoFrameskip-v4-class (A = 6) tree_depth 3, ONE tree, its 216 leaves sharded over the ranks
    by leaf block: every rank expands the first levels, keeps its block and resumes from there (stm_bfs_resume), evaluates its
    leaves with K4, reduces them per first action and the ranks exchange 2*A floats (NCCL all-gather) to finish the logits.
    Each step is one SoftTreeMax policy evaluation of the root; the result is checked against the whole tree on one GPU."""
    import torch
    import torch.distributed as dist
    from softtreemax_b200 import shard
    from softtreemax_b200.engine import Engine
    from softtreemax_b200.roots import make_roots
    game, depth = "synth_pong", 3
    own = eng is None
    if own:
        eng = Engine(game, device=local)
    A = eng.n_actions
    states, stacks = make_roots(eng, 1, seed=1234)                       # the same root on every rank
    eng.cnn_set_weights(_sb3_params(A))
    val_coef, beta = 0.99 ** depth, 1.0

    def one():
        blk = shard.expand_leaf_block(eng, states[0], stacks[0], depth, rank, world)
        if blk["stacks"].shape[0]:
            head, _ = eng.cnn_forward(blk["stacks"].contiguous())
        else:
            head = torch.empty((0, A), device=eng.device)
        stat, cnt = shard.partial_action_stats(head, blk["rewards"], blk["first_action"], A, val_coef, beta)
        return shard.combine_action_stats(stat, cnt)

    for _ in range(3):
        raw = one()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        raw = one()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = shard.max_over_ranks([e0.elapsed_time(e1)], eng.device)[0]
    # whole tree on this GPU, for the check
    full = eng.bfs(states, stacks, depth)
    head, _ = eng.cnn_forward(full["stacks"])
    fa = torch.arange(A ** depth, device=eng.device) // (A ** (depth - 1))
    st, ct = shard.partial_action_stats(head, full["rewards"], fa, A, val_coef, beta)
    ref = st - torch.log(ct)
    err = float((raw - ref).abs().max())
    if own:
        eng.close()
    return {"metric": "softtreemax_policy_evals_per_sec", "mode": "leafshard", "value": steps / (ms / 1e3), "unit": "evals/s",
            "ms_per_step": ms / steps, "n_gpus": world, "steps": steps,
            "leaves_per_sec": A ** depth * steps / (ms / 1e3),
            "config": {"workload": f"{game} (stands for PongNoFrameskip-v4, A={A}) tree_depth={depth}, one tree of {A ** depth} leaves "
                                   f"split into {world} leaf block(s)", "scaling": "strong"},
            "collective": {"op": "all_gather", "backend": "nccl" if world > 1 else "none (one rank)", "bytes_per_rank": 2 * A * 4,
                           "what": "per-first-action partial logsumexp + count of the leaf logits (actor_critic_ts.py:84-85)"},
            "max_abs_diff_vs_one_gpu_tree": err}


def run_train(args, rank, world, local, steps):
    """One PPO minibatch per step on every rank (main.py:54: batch 32, depth-2 Breakout-class tree, 16 leaves per root):
    policy.evaluate_actions on 32 cached roots (K4 on 32*L + 32 images forward and backward, K5 forward and backward),
    clipped-surrogate loss, backward, gradient all-reduce over the ranks (one flat fp32 bucket, NCCL), Adam step."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from softtreemax_b200 import shard
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    dev = torch.device("cuda", local)
    kw = dict(env_name="BreakoutNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    np.random.seed(100 + rank)
    torch.manual_seed(4)
    env = CuleEnv(device=dev, env_kwargs=kw, clip_reward=True, fire_reset=True)
    depth, B = 2, 32
    pol = ActorCriticCnnTSPolicy(env.observation_space, env.action_space, lambda _: 2.5e-4, tree_depth=depth, gamma=0.99, step_env=env,
                                 buffer_size=128, learn_alpha=False, learn_beta=True, max_width=-1, use_leaves_v=False,
                                 is_cumulative_mode=False, regularization=0.001, device=dev)
    shard.broadcast_parameters(pol.parameters())
    obs = env.reset()
    obs_l, act_l, logp_l = [], [], []
    while len(obs_l) < B:
        with torch.no_grad():
            a, v, lp = pol.forward(torch.from_numpy(np.asarray(obs)).float()[None])
        key = int(pol.hash_obs(torch.from_numpy(np.asarray(obs)).float()[None])[0])
        if all(key != k for k, *_ in obs_l):
            obs_l.append((key, np.asarray(obs).copy())); act_l.append(int(a)); logp_l.append(float(lp))
        obs, _, done, _ = env.step(int(a))
        if done:
            obs = env.reset()
    ob = torch.from_numpy(np.stack([o for _, o in obs_l])).float()
    act = torch.tensor(act_l, device=dev)
    old_lp = torch.tensor(logp_l, device=dev)
    adv = torch.linspace(-1, 1, B, device=dev)
    L = env.action_space.n ** depth

    def one():
        values, logp, ent = pol.evaluate_actions(ob, act)
        ratio = torch.exp(logp - old_lp)
        pg = -torch.min(adv * ratio, adv * torch.clamp(ratio, 0.9, 1.1)).mean()
        loss = pg + 0.01 * (-ent.mean()) + 1.0 * torch.nn.functional.mse_loss(torch.zeros(B, device=dev), values.reshape(-1))
        pol.optimizer.zero_grad()
        loss.backward()
        shard.allreduce_gradients(pol.parameters())
        torch.nn.utils.clip_grad_norm_(pol.parameters(), 0.5)
        pol.optimizer.step()
        return loss

    for _ in range(2):
        one()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = one()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = shard.max_over_ranks([e0.elapsed_time(e1)], dev)[0]
    n_par = sum(p.numel() for p in pol.parameters() if p.requires_grad)
    # tensor-core roofline of the update: forward + data gradients + weight gradients ~ 3 x the forward FLOPs of every image
    # (conv1 has no data gradient: 2 x), against the measured bf16 peak
    A_ = env.action_space.n
    fwd = 18685952 + 1024 * (A_ + 1)
    flop_per_image = 3 * fwd - 2 * 400 * 256 * 32
    peak_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    tf_peak = float(json.load(open(peak_path))["bf16_tflops"]) if os.path.exists(peak_path) else 1590.0
    tflops = (B * L + B) * flop_per_image * steps / (ms / 1e3) / 1e12
    return {"metric": "ppo_update_images_per_sec", "mode": "train", "value": world * (B * L + B) * steps / (ms / 1e3), "unit": "images/s",
            "roofline": {"bound": "tensor", "achieved": tflops, "peak": tf_peak, "unit": "TFLOP/s", "frac": tflops / tf_peak,
                         "note": "per rank; a 544-image minibatch is ~45 small launches: launch latency, not the tensor pipe, sets the time"},
            "ms_per_step": ms / steps, "n_gpus": world, "steps": steps, "loss": float(loss.detach()),
            "config": {"workload": f"synth_breakout tree_depth={depth}: one PPO minibatch of {B} cached roots per rank = {B * L + B} images "
                                   "forward + backward through NatureCNN, K5 forward + backward, Adam", "scaling": "weak"},
            "collective": {"op": "all_reduce(sum) of the gradient", "backend": "nccl" if world > 1 else "none (one rank)",
                           "bytes": 4 * n_par, "parameters": n_par}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--roots", type=int, default=4096, help="root states per GPU")
    ap.add_argument("--depth", type=int, default=2)
    ap.add_argument("--seed", type=int, default=1234, help="seed of the first synthetic game the roots are taken from")
    ap.add_argument("--games", type=int, default=8,
                    help="the roots of every rank are taken, 192 at a time, from this many games (seeds seed .. seed + games - 1), "
                         "rank r starting at game r: every rank expands the same mix")
    ap.add_argument("--game", default=GAME, help="cartridge name (softtreemax_b200/roms/games.json)")
    ap.add_argument("--no-policy", action="store_true", help="skip the bfs + CNN + logits arm (sweeps)")
    ap.add_argument("--two-groups", action="store_true", help="time two root groups in flight on two streams even with --no-policy")
    ap.add_argument("--no-interp", action="store_true", help="skip the interpreter-only figure (sweeps)")
    ap.add_argument("--no-collectives", action="store_true", help="skip the leaf-shard and PPO-minibatch arms (the two NCCL collectives)")
    ap.add_argument("--mode", default="bfs", choices=["bfs", "leafshard", "train"],
                    help="bfs: the headline (plus every side arm); leafshard / train: only that arm (BASELINE.json configs[2] / the PPO update)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--scratch-envs", type=int, default=262144, help="envs per K2/K3 pass (native-frame scratch: 67 KB each)")
    ap.add_argument("--profile", action="store_true",
                    help="bracket one bfs step + one policy step with cudaProfilerStart/Stop and exit "
                         "(for `ncu --profile-from-start off`); prints no bench line")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from softtreemax_b200.engine import Engine
    from softtreemax_b200.roots import make_roots

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL's own log (communicator size, NVLS / ring choice) is wanted, but it goes to file descriptor 1: keep a private
        # copy of stdout for the one JSON line and point descriptor 1 at stderr for everything else.
        # (NCCL_DEBUG is left as the launcher set it)
        sys.stdout.flush()
        global _JSON_OUT
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        chk = torch.tensor([float(rank), 1.0], device=torch.device("cuda", local))
        dist.all_reduce(chk)                                     # every rank took part: sum of ranks and count of ranks
        nccl_check = {"backend": "nccl", "ranks_counted": int(chk[1].item()), "sum_of_ranks_ok": int(chk[0].item()) == world * (world - 1) // 2,
                      "nccl_version": ".".join(str(v) for v in torch.cuda.nccl.version())}
    if world == 1:
        nccl_check = {"backend": "none (one rank)", "ranks_counted": 1, "sum_of_ranks_ok": True}
    if args.mode != "bfs":
        res = (run_leafshard(args, None, rank, world, local, steps=args.steps) if args.mode == "leafshard"
               else run_train(args, rank, world, local, steps=args.steps))
        if rank == 0:
            print(json.dumps(res), file=_JSON_OUT or sys.stdout, flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    A, depth, R = n_actions_of(args.game), args.depth, args.roots
    L = R * A ** depth
    eng = Engine(args.game, device=local, scratch_envs=min(L, args.scratch_envs))
    dev = eng.device
    # every rank owns its own shard of roots (different seeds => different states per rank)
    states, stacks = make_roots(eng, R, seed=args.seed, trajectories=args.games, first=rank)
    out = dict(stacks=torch.empty((L, 4, 84, 84), dtype=torch.uint8, device=dev),
               rewards=torch.empty((L,), dtype=torch.float32, device=dev),
               done=torch.empty((L,), dtype=torch.uint8, device=dev),
               lives=torch.empty((L,), dtype=torch.int32, device=dev), states=None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident arm ----------------
    for _ in range(args.warmup):
        eng.bfs(states, stacks, depth, out=out)
    barrier()
    if args.profile:
        from softtreemax_b200 import ops as _ops
        eng.cnn_set_weights({k: torch.randn(s) * 0.05 for k, s in {
            "cnn.0.weight": (32, 4, 8, 8), "cnn.0.bias": (32,), "cnn.2.weight": (64, 32, 4, 4), "cnn.2.bias": (64,),
            "cnn.4.weight": (64, 64, 3, 3), "cnn.4.bias": (64,), "linear.0.weight": (512, 3136), "linear.0.bias": (512,),
            "action_net.weight": (A, 512), "action_net.bias": (A,), "value_net.weight": (1, 512), "value_net.bias": (1,)}.items()})
        eng.cnn_forward(stacks)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        eng.bfs(states, stacks, depth, out=out)
        hl, _ = eng.cnn_forward(out["stacks"])
        _ops.softtreemax_logits_full(eng, hl.reshape(R, A ** depth, A), out["rewards"].reshape(R, -1), 0.98, 1.0, 1.0, 0.001, False)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        eng.close()
        return
    eng.timing(True)
    launches0 = eng.launch_count
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        eng.bfs(states, stacks, depth, out=out)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop()
    launches = eng.launch_count - launches0
    per_kernel = eng.timing_read()
    eng.timing(False)

    # ---------------- interpreter only (a cartridge image the library has no translation for) ----------------
    # The headline runs the block-translated kernel of this very cartridge image (tools/xlate_rom.cpp, matched by hash).
    # Any other image — a user-supplied ROM — gets the interpreter: the same step is timed with STM_XLATE=0 beside it.
    ms_interp = 0.0
    if not args.no_interp:
        os.environ["STM_XLATE"] = "0"
        eng_i = Engine(args.game, device=local, scratch_envs=min(L, args.scratch_envs))
        del os.environ["STM_XLATE"]
        for _ in range(2):
            eng_i.bfs(states, stacks, depth, out=out)
        barrier()
        i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_i = max(2, args.steps // 2)
        i0.record()
        for _ in range(n_i):
            eng_i.bfs(states, stacks, depth, out=out)
        i1.record()
        barrier()
        ms_interp = i0.elapsed_time(i1) / n_i
        eng_i.close()

    # ---------------- two root groups in flight (throughput mode) ----------------
    # A depth-d step is d dependent K2 launches, each bound by the latency of one emulated CPU per lane, and none fills the
    # issue slots.  Two independent groups of roots (two contexts, two streams), started half a step apart, overlap one
    # group's small first level with the other's last level — what a rollout with two env groups taking turns does.
    # Reported beside the headline, which stays one group at a time.
    ms_two = 0.0
    if args.two_groups or not args.no_policy:
        eng_a = Engine(args.game, device=local, scratch_envs=min(L, args.scratch_envs))
        eng2 = Engine(args.game, device=local, scratch_envs=min(L, args.scratch_envs))
        states2, stacks2 = make_roots(eng2, R, seed=args.seed, trajectories=args.games, first=rank + args.games // 2)
        out_b = dict(stacks=torch.empty_like(out["stacks"]), rewards=torch.empty_like(out["rewards"]),
                     done=torch.empty_like(out["done"]), lives=torch.empty_like(out["lives"]), states=None)
        s_a, s_b = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

        def two_groups(nsteps):
            main = torch.cuda.current_stream()
            s_a.wait_stream(main)
            s_b.wait_stream(main)
            with torch.cuda.stream(s_b):
                torch.cuda._sleep(int(0.45 * ms_total / args.steps * 1e-3 * 1.9e9))     # start group B about half a step later
            for _ in range(nsteps):
                with torch.cuda.stream(s_a):
                    eng_a.bfs(states, stacks, depth, out=out)
                with torch.cuda.stream(s_b):
                    eng2.bfs(states2, stacks2, depth, out=out_b)
            main.wait_stream(s_a)
            main.wait_stream(s_b)

        two_groups(2)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        two_groups(args.steps)
        g1.record()
        barrier()
        ms_two = g0.elapsed_time(g1)
        eng2.close()
        eng_a.close()
        del out_b

    ms_policy = ms_policy_e2e = ms_pipe = c_ms = 0.0
    hs_pin = hk_pin = None
    if not args.no_policy:
        # ---------------- policy arm: bfs + NatureCNN on leaves and roots (K4) + SoftTreeMax logits (K5) ----------
        from softtreemax_b200 import ops
        import math
        gtor = torch.Generator().manual_seed(4)                  # reference default seed, utils.py:67
        shapes = {"cnn.0": (32, 4, 8, 8), "cnn.2": (64, 32, 4, 4), "cnn.4": (64, 64, 3, 3), "linear.0": (512, 3136),
                  "action_net": (A, 512), "value_net": (1, 512)}
        params = {}
        for k, shp in shapes.items():                            # random-init weights of the architecture (orthogonal, SB3 gains)
            w = torch.empty(shp)
            torch.nn.init.orthogonal_(w, gain={"action_net": 0.01, "value_net": 1.0}.get(k, math.sqrt(2)), generator=gtor)
            params[k + ".weight"], params[k + ".bias"] = w, torch.zeros(shp[0])
        eng.cnn_set_weights(params)
        val_coef = 0.99 ** depth

        def policy_step():
            eng.bfs(states, stacks, depth, out=out)
            head_leaf, _ = eng.cnn_forward(out["stacks"])
            head_root, val_root = eng.cnn_forward(stacks)
            logits, _, _ = ops.softtreemax_logits_full(eng, head_leaf.reshape(R, A ** depth, A), out["rewards"].reshape(R, -1),
                                                      val_coef, 1.0, 1.0, 0.001, False)
            return logits, val_root

        for _ in range(2):
            policy_step()
        barrier()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c_ms = 0.0
        p0.record()
        for _ in range(args.steps):
            policy_step()
        p1.record()
        barrier()
        ms_policy = p0.elapsed_time(p1)
        # pipelined: tree expansion of batch i+1 (stream A) overlaps CNN + logits of batch i (stream B);
        # two leaf buffers, events for the hand-offs.  K2 leaves issue slots and the tensor pipe idle,
        # K4 leaves the ALUs idle: they co-run.
        out2 = dict(stacks=torch.empty_like(out["stacks"]), rewards=torch.empty_like(out["rewards"]),
                    done=torch.empty_like(out["done"]), lives=torch.empty_like(out["lives"]), states=None)
        bufs = [out, out2]
        s_tree, s_net = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        free = [torch.cuda.Event(), torch.cuda.Event()]

        def pipelined(nsteps):
            main = torch.cuda.current_stream()
            s_tree.wait_stream(main)
            s_net.wait_stream(main)
            for i in range(nsteps):
                b = i & 1
                with torch.cuda.stream(s_tree):
                    if i >= 2:
                        s_tree.wait_event(free[b])
                    eng.bfs(states, stacks, depth, out=bufs[b])
                    ready[b].record(s_tree)
                with torch.cuda.stream(s_net):
                    s_net.wait_event(ready[b])
                    hl, _ = eng.cnn_forward(bufs[b]["stacks"])
                    eng.cnn_forward(stacks)
                    ops.softtreemax_logits_full(eng, hl.reshape(R, A ** depth, A), bufs[b]["rewards"].reshape(R, -1),
                                                val_coef, 1.0, 1.0, 0.001, False)
                    free[b].record(s_net)
            main.wait_stream(s_tree)
            main.wait_stream(s_net)

        pipelined(2)
        barrier()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        pipelined(args.steps)
        r1.record()
        barrier()
        ms_pipe = r0.elapsed_time(r1)
        # K4 alone, same inputs, for its tensor roofline
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(args.steps):
            eng.cnn_forward(out["stacks"])
        c1.record()
        torch.cuda.synchronize()
        c_ms = c0.elapsed_time(c1)
        # policy end to end: host roots in, host logits + values out
        h_logits = torch.empty((R, A), dtype=torch.float32).pin_memory()
        h_values = torch.empty((R,), dtype=torch.float32).pin_memory()
        hs_pin, hk_pin = states.cpu().pin_memory(), stacks.cpu().pin_memory()

        s_up = torch.cuda.Stream(device=dev)
        up_done = torch.cuda.Event()

        def policy_step_host():
            # the states (1 MB) on the compute stream, the stacks (116 MB) on their own: the tree starts at once and only its
            # first stack shift — and the CNN on the root observations — wait for them (stm_bfs_after)
            main = torch.cuda.current_stream()
            s_up.wait_stream(main)                          # the previous step is done with `stacks`
            states.copy_(hs_pin, non_blocking=True)
            with torch.cuda.stream(s_up):
                stacks.copy_(hk_pin, non_blocking=True)
                up_done.record(s_up)
            eng.bfs(states, stacks, depth, out=out, stacks_ready=up_done)
            head_leaf, _ = eng.cnn_forward(out["stacks"])
            main.wait_event(up_done)
            head_root, vr = eng.cnn_forward(stacks)
            lg, _, _ = ops.softtreemax_logits_full(eng, head_leaf.reshape(R, A ** depth, A), out["rewards"].reshape(R, -1),
                                                  val_coef, 1.0, 1.0, 0.001, False)
            h_logits.copy_(lg, non_blocking=True)
            h_values.copy_(vr, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        policy_step_host()
        barrier()
        q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        q0.record()
        for _ in range(args.steps):
            policy_step_host()
        q1.record()
        barrier()
        ms_policy_e2e = q0.elapsed_time(q1)

    # ---------------- end-to-end arm: host roots in, host rewards/done/lives out ----------------
    h_states = states.cpu().pin_memory()
    h_stacks = stacks.cpu().pin_memory()
    h_out = dict(stacks=out["stacks"], rewards=torch.empty((L,), dtype=torch.float32).pin_memory(),
                 done=torch.empty((L,), dtype=torch.uint8).pin_memory(),
                 lives=torch.empty((L,), dtype=torch.int32).pin_memory())
    np_out = dict(stacks=h_out["stacks"], rewards=h_out["rewards"].numpy(), done=h_out["done"].numpy(), lives=h_out["lives"].numpy())
    for _ in range(2):
        eng.bfs_host(h_states.numpy(), h_stacks.numpy(), depth, out=np_out)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        eng.bfs_host(h_states.numpy(), h_stacks.numpy(), depth, out=np_out)
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    h2d = int(h_states.numel() + h_stacks.numel())
    d2h = int(L * (4 + 1 + 4))

    from softtreemax_b200 import shard
    # every rank's own step time and kernel times (each rank expands different roots: the step time is the max over ranks)
    mine = torch.tensor([ms_total / args.steps, per_kernel["step"][0] / args.steps, per_kernel["frames"][0] / args.steps,
                         per_kernel["clone"][0] / args.steps, ms_interp], dtype=torch.float64, device=dev)
    if world > 1:
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
    else:
        allr = [mine]
    per_rank = [dict(rank=i, ms_per_step=float(t[0]), k2_ms=float(t[1]), k3_ms=float(t[2]), k1_ms=float(t[3]),
                     interpreter_only_ms_per_step=float(t[4])) for i, t in enumerate(allr)]
    ms_total, ms_e2e, ms_policy, ms_policy_e2e, ms_pipe, ms_two, ms_interp = shard.max_over_ranks(
        [ms_total, ms_e2e, ms_policy, ms_policy_e2e, ms_pipe, ms_two, ms_interp], dev)
    collectives = None
    if not args.no_collectives:
        collectives = {"leaf_shard": run_leafshard(args, eng if args.game == "synth_pong" else None, rank, world, local, steps=max(3, args.steps)),
                       "ppo_minibatch": run_train(args, rank, world, local, steps=max(3, min(args.steps, 6)))}

    if rank == 0:
        ms_step = ms_total / args.steps
        value = shard.whole_job_rate(L * args.steps, world, ms_total)
        e2e = shard.whole_job_rate(L * args.steps, world, ms_e2e)
        hbm_peak, peak_src = peaks()
        S = 256
        step_ms, step_n = per_kernel["step"]
        clone_ms, clone_n = per_kernel["clone"]
        frames_ms, frames_n = per_kernel["frames"]
        shift_ms, shift_n = per_kernel["shift"]
        node_steps = R * sum(A ** k for k in range(1, depth + 1))            # per bfs call
        leaves_like = node_steps                                              # envs K2/K3 touched per bfs call
        # algorithmic bytes per bfs call
        k2_bytes = node_steps * (2 * S + 2 * (4 + 1 + 4))
        k1_bytes = sum(R * A ** k * S + R * A ** (k + 1) * S for k in range(depth))
        k3_bytes = node_steps * 49392
        def gbs(b, ms):
            return b * args.steps / (ms / 1e3) / 1e9 if ms > 0 else None
        k2 = gbs(k2_bytes, step_ms)
        # K2 is bound by issue slots / the latency of one emulated CPU per lane, not by HBM: its roofline is the
        # warp-instruction issue rate.  achieved = warp instructions per node-step (ncu, profiles/) x node-steps / K2 time;
        # peak = one warp instruction per cycle on each of the 148 x 4 schedulers at the SM clock sampled during the run.
        sm_hz = 1e6 * float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965)
        issue_peak = 148 * 4 * sm_hz / 1e9
        issue = K2_NCU_WARP_INST_PER_ENV * node_steps * args.steps / (step_ms / 1e3) / 1e9
        roofline = {"kernel": "stm_step_kernel (K2: 6507+TIA+RIOT agent step)", "bound": "issue",
                    "achieved": issue, "peak": issue_peak, "unit": "G warp-instructions/s", "frac": issue / issue_peak,
                    "achieved_source": "smsp__inst_executed.sum per env from the ncu capture (profiles/r02_k2_final_ncu_full_summary.txt: "
                                       "%.0f warp instructions per node-step) x node-steps / the K2 time measured here" % K2_NCU_WARP_INST_PER_ENV,
                    "peak_source": "148 SMs x 4 schedulers x SM clock sampled during the timed region",
                    "issue_active_ncu": K2_NCU_ISSUE_ACTIVE,
                    "traffic": K2_NCU_DRAM_BYTES_PER_ENV * node_steps / max(1, step_n / args.steps),
                    "traffic_source": "ncu --set full, profiles/r02_k2_final_ncu_full_summary.txt (per env) x envs per launch",
                    "avg_launch_ms": step_ms / max(1, step_n), "launches": step_n,
                    "share_of_step": step_ms / ms_total,
                    "timing": "CUDA event pair around every K1/K2/K3 launch on the launch stream (stm_timing_*), inside the timed loop",
                    "hbm": {"bound": "hbm", "achieved": k2, "peak": hbm_peak, "unit": "GB/s", "frac": k2 / hbm_peak if k2 else None,
                            "algorithmic_bytes_per_launch": k2_bytes / max(1, step_n / args.steps), "peak_source": peak_src,
                            "note": "HBM floor 2*S B/node-step: small by construction, the kernel is not memory bound"},
                    "emulated_cycles_per_s": node_steps * CYCLES_PER_NODE_STEP * args.steps / (step_ms / 1e3),
                    "node_steps_per_s": node_steps * args.steps / (step_ms / 1e3)}
        kernels = {
            "K1_clone": {"bound": "hbm", "achieved": gbs(k1_bytes, clone_ms), "peak": hbm_peak, "unit": "GB/s",
                         "frac": (gbs(k1_bytes, clone_ms) or 0) / hbm_peak, "ms_per_step": clone_ms / args.steps,
                         "launches": clone_n, "share_of_step": clone_ms / ms_total},
            # K3 = the newest plane (stm_frames_kernel, caller's stream) + the stack shift (K3b, stm_shift_kernel, on stm_bfs's side
            # stream under the K2 launch of the same level).  The fraction is taken on the SUM of the two kernels' durations.
            "K3_frames": {"bound": "hbm", "achieved": gbs(k3_bytes, frames_ms + shift_ms), "peak": hbm_peak, "unit": "GB/s",
                          "frac": (gbs(k3_bytes, frames_ms + shift_ms) or 0) / hbm_peak, "ms_per_step": frames_ms / args.steps,
                          "launches": frames_n, "share_of_step": frames_ms / ms_total,
                          "algorithmic_bytes_per_leaf": 49392,
                          "shift_ms_per_step": shift_ms / args.steps, "shift_launches": shift_n,
                          "shift_GBps": gbs(node_steps * 21168 * (1 + 1 / A), shift_ms),
                          "note": "frac = 49 392 B/leaf over (frames + shift) kernel time; the shift runs concurrently with K2, "
                                  "so only ms_per_step (frames) is on the step's critical path"},
            "K2_step": {"ms_per_step": step_ms / args.steps, "launches": step_n, "share_of_step": step_ms / ms_total},
        }
        tf_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) \
            if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 1590.0
        cnn_tflops = L * args.steps * (18685952 + 1024 * (A + 1)) / (c_ms / 1e3) / 1e12 if c_ms > 0 else 0.0
        if not args.no_policy:
            kernels["K4_cnn"] = {"bound": "tensor", "achieved": cnn_tflops, "peak": tf_peak, "unit": "TFLOP/s",
                                 "frac": cnn_tflops / tf_peak, "ms_per_step": c_ms / args.steps,
                                 "images_per_step": L, "dtype": "bf16 x bf16 -> f32 (tcgen05)"}
        policy = None if args.no_policy else {"metric": "softtreemax_policy_evals_per_sec", "value": world * R * args.steps / (ms_policy / 1e3),
                  "unit": "evals/s", "ms_per_step": ms_policy / args.steps,
                  "leaves_per_sec_through_cnn": world * L * args.steps / (ms_policy / 1e3),
                  "pipelined": {"value": world * R * args.steps / (ms_pipe / 1e3), "unit": "evals/s",
                                "ms_per_step": ms_pipe / args.steps,
                                "how": "two streams: bfs of batch i+1 overlaps CNN + logits of batch i (double-buffered leaves)"},
                  "e2e": {"value": world * R * args.steps / (ms_policy_e2e / 1e3), "unit": "evals/s",
                          "h2d_bytes_per_step": int(hs_pin.numel() + hk_pin.numel()), "d2h_bytes_per_step": int(R * (A + 1) * 4),
                          "api": "pinned host roots -> stm_bfs_after (stacks uploaded on their own stream) + stm_cnn_forward + stm_logits_fwd -> host logits, values"},
                  "what": "bfs + NatureCNN (K4) on R*A^d leaves and R roots + SoftTreeMax logits (K5); random-init SB3-style weights"}
        cpu_rate, cpu_leaves, cpu_dt = cpu_oracle_rate(depth, args.cpu_seconds, os.cpu_count() or 1, game=args.game)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": workload_config(args),
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": ms_e2e / args.steps,
                        "api": "stm_bfs_host: pinned host roots in; rewards/done/lives to host; leaf stacks stay in HBM"},
                "two_groups_in_flight": None if not ms_two else {
                    "value": world * 2 * L * args.steps / (ms_two / 1e3), "unit": UNIT, "ms_per_pair_of_steps": ms_two / args.steps,
                    "how": "two independent groups of roots_per_gpu roots, two contexts on two streams half a step apart"},
                "interpreter_only": None if not ms_interp else {
                    "value": world * L / (ms_interp / 1e3), "unit": UNIT, "ms_per_step": ms_interp,
                    "what": "the same step with STM_XLATE=0: what a cartridge image gets that has neither a bundled translation nor a plug-in (translate.py)"},
                "translated_core": int(eng.translated_core),
                "per_rank": per_rank, "collectives": collectives, "comm": nccl_check,
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "kernels": kernels, "policy": policy,
                "cpu_baseline": {"value": cpu_rate, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                 "sample": f"{cpu_leaves} leaves in {cpu_dt:.1f}s: 64 roots x depth {depth} repeated, oracle port (OpenMP)"}}
        print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
