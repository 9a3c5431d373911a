"""Throughput sweep over root counts, tree depths and action-set sizes (BASELINE.json configs[4]).

Runs bench.py once per cell (tree expansion only: --no-policy) and writes every JSON line plus a
markdown table.  Cells are bounded so that the leaf stacks of one step stay below ~32 GB.  Action-set sizes cover the
reference's sweep (yamls/example_sweep.yml:19-26: 14, 4, 8, 18, 14, 6, 8, 9), depth 0 is one agent step of the roots without a
tree (the many-env arm), R = 1 is the reference's own single-root shape.

    python tools/sweep.py gpurun_out/sweep.json [--quick]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CELLS = [("synth_breakout", r, d) for r in (256, 1024, 4096) for d in (0, 1, 2, 3)] + \
        [("synth_breakout", 256, 4), ("synth_breakout", 1024, 4)] + \
        [("synth_pong", r, d) for r in (256, 4096) for d in (0, 1, 2, 3)] + \
        [("synth_pacman", 256, 2), ("synth_pacman", 4096, 2), ("synth_pacman", 256, 3)] + \
        [("synth_joy8", 256, 2), ("synth_joy8", 4096, 2), ("synth_joy8", 256, 3)] + \
        [("synth_joy14", 256, 2), ("synth_joy14", 2048, 2), ("synth_joy14", 256, 3)] + \
        [("synth_joy18", 256, 0), ("synth_joy18", 256, 1), ("synth_joy18", 256, 2), ("synth_joy18", 1024, 2), ("synth_joy18", 128, 3)] + \
        [("synth_torture", 256, 1), ("synth_torture", 4096, 1), ("synth_torture", 256, 2)] + \
        [("synth_breakout", 1, 2), ("synth_breakout", 1, 3), ("synth_pong", 1, 3)]          # R = 1: the reference's own shape (main.py:47)
QUICK = [("synth_breakout", 256, 0), ("synth_breakout", 256, 1), ("synth_pong", 256, 2), ("synth_breakout", 1, 2)]


def depth0_cell(game, roots, steps=20):
    """tree_depth = 0: no tree — one agent step of `roots` root envs (K2 + K3), the many-env arm of main.py:42-45.
    leaves = roots; timed in this process with CUDA events."""
    import torch
    sys.path.insert(0, ROOT)
    from softtreemax_b200.engine import Engine
    from softtreemax_b200.roots import make_roots
    eng = Engine(game, device=0, scratch_envs=max(roots, 4096))
    states, stacks = make_roots(eng, roots, seed=1234, trajectories=8)
    rw = torch.zeros(roots, device=eng.device)
    dn = torch.zeros(roots, dtype=torch.uint8, device=eng.device)
    lv = torch.full((roots,), 9, dtype=torch.int32, device=eng.device)
    act = (torch.arange(roots, device=eng.device) % eng.n_actions).to(torch.uint8)
    out = torch.empty_like(stacks)

    def step():
        dn.zero_()
        eng.step(states, rw, dn, lv, actions=act, frameskip=4, episodic_life=False, weight=1.0, render=True)
        eng.frames(stacks, roots, fanout=1, out=out)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    A = eng.n_actions
    eng.close()
    return {"value": roots / (ms / 1e3), "ms_per_step": ms, "config": {"leaves_per_gpu": roots, "roots_per_gpu": roots},
            "e2e": {"value": float("nan")}, "roofline": {"node_steps_per_s": roots / (ms / 1e3)}, "cpu_baseline": {"value": float("nan")},
            "n_actions": A}


def main():
    out = sys.argv[1]
    cells = QUICK if "--quick" in sys.argv else CELLS
    lines = []
    for game, roots, depth in cells:
        if depth == 0:
            d = depth0_cell(game, roots)
            lines.append({"game": game, "roots": roots, "depth": depth, "line": d})
            print(game, roots, depth, round(d["value"]), flush=True)
            continue
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--game", game, "--roots", str(roots), "--depth", str(depth),
               "--steps", "3", "--warmup", "3", "--no-policy", "--no-interp", "--no-collectives", "--cpu-seconds", "0.3"]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        js = [l for l in r.stdout.splitlines() if l.startswith("{")]
        if r.returncode or not js:
            lines.append({"game": game, "roots": roots, "depth": depth, "error": (r.stderr or r.stdout)[-400:]})
            continue
        d = json.loads(js[-1])
        lines.append({"game": game, "roots": roots, "depth": depth, "line": d})
        print(game, roots, depth, round(d["value"]), flush=True)
    json.dump(lines, open(out, "w"), indent=1)
    md = ["| cartridge | A | roots | depth | leaves / step | ms / step | leaves / s | e2e leaves / s | node-steps / s | CPU oracle leaves / s |",
          "|---|---|---|---|---|---|---|---|---|---|"]
    for c in lines:
        if "error" in c:
            md.append(f"| {c['game']} | | {c['roots']} | {c['depth']} | error: {c['error'][-80:]} | | | | | |")
            continue
        d = c["line"]
        cfg = d["config"]
        A = d["n_actions"] if c["depth"] == 0 else round(cfg["leaves_per_gpu"] / cfg["roots_per_gpu"]) ** (1.0 / c["depth"])
        md.append(f"| {c['game']} | {round(A)} | {c['roots']} | {c['depth']} | {cfg['leaves_per_gpu']} | {d['ms_per_step']:.2f} | "
                  f"{d['value'] / 1e6:.2f} M | {d['e2e']['value'] / 1e6:.2f} M | {d['roofline']['node_steps_per_s'] / 1e6:.2f} M | "
                  f"{d['cpu_baseline']['value'] / 1e3:.1f} k |")
    open(os.path.splitext(out)[0] + ".md", "w").write("\n".join(md) + "\n")


if __name__ == "__main__":
    main()
