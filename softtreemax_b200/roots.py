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


def make_roots(eng, R, seed=1234, base_steps=64, spread=192):
    rng = np.random.default_rng(seed)
    acts = rng.integers(0, eng.n_actions, size=base_steps + spread)
    st, rw, dn, lv = eng.reset_states(1)
    stack = torch.zeros((1, 4, 84, 84), dtype=torch.uint8, device=eng.device)
    need = sorted(set(base_steps + (i % spread) for i in range(R)))
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
    idx = [base_steps + (i % spread) for i in range(R)]
    states = torch.cat([snaps[n][0] for n in idx]).contiguous()
    stacks = torch.cat([snaps[n][1] for n in idx]).contiguous()
    return states, stacks
