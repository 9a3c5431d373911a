"""Generates tests/golden/oracle_traces.json: hashes of oracle state / observation / bfs output on
seeded action sequences for every cartridge.  These freeze the oracle's behaviour (a change to the
emulator model must be deliberate and regenerate them).  Self-authored: the reference ships no
vectors for this path.

    python tools/gen_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def h(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def trace(orc, name, steps=120):
    g = orc.Game(name)
    rng = np.random.default_rng(7)
    s = g.reset()
    done = np.zeros(1, np.uint8)
    rew = np.zeros(1, np.float32)
    lives = np.array([g.lives(s)], np.int32)
    obs_h = hashlib.sha1()
    for t in range(steps):
        done[:] = 0
        obs, *_ = g.agent_step(s, int(rng.integers(0, g.n_actions)), done=done, reward=rew, lives=lives)
        obs_h.update(obs.tobytes())
    roots, stacks = orc.make_roots(g, 3, seed=11, base_steps=20, spread=7)
    out = g.bfs(roots, stacks, 2 if g.n_actions <= 6 else 1)
    return dict(state=h(s), obs=obs_h.hexdigest()[:16], reward=float(rew[0]), lives=int(lives[0]),
                bfs_stack=h(out[0]), bfs_reward=h(out[1]), bfs_done=h(out[2]), bfs_state=h(out[4]))


if __name__ == "__main__":
    from oracle import orc
    games = json.load(open(os.path.join(ROOT, "softtreemax_b200", "roms", "games.json")))
    gold = {name: trace(orc, name) for name in sorted(games)}
    json.dump(gold, open(os.path.join(ROOT, "tests", "golden", "oracle_traces.json"), "w"), indent=1, sort_keys=True)
    print(json.dumps(gold, indent=1))
