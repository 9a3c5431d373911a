"""Throughput sweep over root counts, tree depths and action-set sizes (BASELINE.json configs[4]).

Runs bench.py once per cell (tree expansion only: --no-policy) and writes every JSON line plus a
markdown table.  Cells are bounded so that the leaf stacks of one step stay below ~32 GB.

    python tools/sweep.py gpurun_out/sweep.json [--quick]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CELLS = [("synth_breakout", r, d) for r in (256, 1024, 4096) for d in (1, 2, 3)] + \
        [("synth_breakout", 256, 4), ("synth_breakout", 1024, 4)] + \
        [("synth_pong", r, d) for r in (256, 4096) for d in (2, 3)] + \
        [("synth_pacman", 256, 2), ("synth_pacman", 4096, 2), ("synth_pacman", 256, 3)]
QUICK = [("synth_breakout", 256, 1), ("synth_pong", 256, 2)]


def main():
    out = sys.argv[1]
    cells = QUICK if "--quick" in sys.argv else CELLS
    lines = []
    for game, roots, depth in cells:
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--game", game, "--roots", str(roots), "--depth", str(depth),
               "--steps", "3", "--warmup", "3", "--no-policy", "--cpu-seconds", "0.3"]
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
        A = round(cfg["leaves_per_gpu"] / cfg["roots_per_gpu"]) ** (1.0 / c["depth"])
        md.append(f"| {c['game']} | {round(A)} | {c['roots']} | {c['depth']} | {cfg['leaves_per_gpu']} | {d['ms_per_step']:.2f} | "
                  f"{d['value'] / 1e6:.2f} M | {d['e2e']['value'] / 1e6:.2f} M | {d['roofline']['node_steps_per_s'] / 1e6:.2f} M | "
                  f"{d['cpu_baseline']['value'] / 1e3:.1f} k |")
    open(os.path.splitext(out)[0] + ".md", "w").write("\n".join(md) + "\n")


if __name__ == "__main__":
    main()
