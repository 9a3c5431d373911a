"""stm_step_frames_noalloc — the entry point for callers that must not make the library allocate
(a host whose own memory pool already owns the GPU; a loop that cannot take cudaMalloc's device-wide
synchronisation).  It runs the same step + frames kernels against a scratch array compiled into the
library instead of the ctx's cudaMalloc'ed one, so its results must be identical to stm_step +
stm_frames: bit-exact, the bar of every emulator test here.  Checked on two consecutive calls,
because all calls share that one array and the second call finds the first one's frames in it.

Collected last on purpose (file name): a `-x` run has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def test_noalloc_entry_point_matches_step_plus_frames():
    from softtreemax_b200._lib import check, lib
    from softtreemax_b200.engine import Engine
    eng = Engine("synth_torture")                 # busiest cartridge: every TIA object, mid-line writes
    try:
        n, A = 100, eng.n_actions                 # not a multiple of the 32-env tile on purpose
        dev = eng.device
        g = torch.Generator().manual_seed(3)
        st, rw, dn, lv = eng.reset_states(n)
        acts = torch.randint(0, A, (n,), generator=g, dtype=torch.uint8).to(dev)
        stacks = torch.zeros((n, 4, 84, 84), dtype=torch.uint8, device=dev)
        for _ in range(3):                        # get away from the power-on state, envs diverge by action
            eng.step(st, rw, dn, lv, actions=acts, render=True)
            stacks = eng.frames(stacks, n, fanout=1)
        dn[5] = 1                                 # one frozen env: its new frame must come out all-zero
        ref = [t.clone() for t in (st, rw, dn, lv)]
        new = [t.clone() for t in (st, rw, dn, lv)]
        ref_stack, new_stack = stacks.clone(), stacks.clone()
        for call in range(2):                     # two calls in a row through each path
            acts = torch.randint(0, A, (n,), generator=g, dtype=torch.uint8).to(dev)
            eng.step(*ref, actions=acts, frameskip=4, weight=0.5, render=True)
            ref_stack = eng.frames(ref_stack, n, fanout=1)
            out = torch.empty_like(new_stack)
            check(lib().stm_step_frames_noalloc(eng._h, _p(new[0]), _p(acts), n, 4, 1, 0.5, _p(new[1]), _p(new[2]),
                                                _p(new[3]), _p(new_stack), _p(out), 1, eng._stream()), eng._h)
            new_stack = out
            torch.cuda.synchronize()
            for name, a, b in zip(("states", "rewards", "done", "lives"), ref, new):
                assert torch.equal(a, b), f"{name} differ on call {call}"
            assert torch.equal(ref_stack, new_stack), f"observation stacks differ on call {call}"
            assert int(new_stack[5, 3].max()) == 0 and int(new_stack[6, 3].max()) > 0
        # sizes beyond the built-in scratch are refused, with a message that says why
        big = eng.reset_states(129)
        bs = torch.zeros((129, 4, 84, 84), dtype=torch.uint8, device=dev)
        rc = lib().stm_step_frames_noalloc(eng._h, _p(big[0]), None, 129, 4, 1, 1.0, _p(big[1]), _p(big[2]), _p(big[3]),
                                           _p(bs), _p(bs.clone()), 1, eng._stream())
        assert rc == -1 and b"exceeds the built-in scratch of 128" in lib().stm_last_error(eng._h)
    finally:
        eng.close()
