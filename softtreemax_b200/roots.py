"""Synthetic root states for benchmarks and tests, produced by the CUDA path itself.

BASELINE.md §4: root i = power-on state advanced by base_steps + (i mod spread) agent steps with
actions drawn from numpy.random.default_rng(seed); the root observation stack is the true last four
frames.  This plays the role of the reference's stepping env (environments/cule_env.py:74-93) for
input generation only.
"""
import numpy as np
import torch


def terminal(info, state_row):
    """Game-over predicate of the descriptor, evaluated on one host copy of a state blob."""
    ta = info["term_addr"]
    if ta == 255:
        return False
    return (int(state_row[ta & 0x7F]) & info["term_mask"]) == info["term_value"]


def _trajectory(eng, seed, base_steps, spread, need):
    """One game played from power-on under actions drawn from default_rng(seed): {n: (state, stack) after n agent steps}."""
    rng = np.random.default_rng(seed)
    acts = rng.integers(0, eng.n_actions, size=base_steps + spread)
    st, rw, dn, lv = eng.reset_states(1)
    stack = torch.zeros((1, 4, 84, 84), dtype=torch.uint8, device=eng.device)
    snaps = {}
    t = 0
    for n in need:
        while t < n:
            dn.zero_()
            a = torch.tensor([int(acts[t])], dtype=torch.uint8, device=eng.device)
            eng.step(st, rw, dn, lv, actions=a, render=True)
            stack = eng.frames(stack, 1, fanout=1)
            t += 1
            if terminal(eng.info, st[0].cpu().numpy()):
                st, rw, dn, lv = eng.reset_states(1)
                stack = torch.zeros((1, 4, 84, 84), dtype=torch.uint8, device=eng.device)
        snaps[n] = (st.clone(), stack.clone())
    return snaps


def make_roots(eng, R, seed=1234, base_steps=64, spread=192, trajectories=1, first=0):
    """Root i = the state after base_steps + (i mod spread) agent steps of a game.  With trajectories = T > 1 every block of
    `spread` consecutive roots comes from the game seeded seed + ((first + block) mod T): a multi-GPU run gives rank r
    first = r, so that every rank expands the SAME mix of games in a rotated order — K2's duration depends on what is on
    the screen (5.6 to 7.4 ms per depth-2 step across eight synthetic Breakout games), and ranks that each played ONE game
    of their own would measure that spread, not the machine."""
    need = sorted(set(base_steps + (i % spread) for i in range(R)))
    blocks = (R + spread - 1) // spread
    games = sorted(set((first + b) % trajectories for b in range(blocks)))
    snaps = {t: _trajectory(eng, seed + t, base_steps, spread, need) for t in games}
    pick = [snaps[(first + i // spread) % trajectories][base_steps + (i % spread)] for i in range(R)]
    states = torch.cat([p[0] for p in pick]).contiguous()
    stacks = torch.cat([p[1] for p in pick]).contiguous()
    return states, stacks
