"""The torchcule facade (softtreemax_b200/atari.py) on the GPU: the raw-pointer, frame-by-frame calls the reference's
CuleBFS.bfs makes (policies/cule_bfs.py:139-172) — expand(n), base-class step(bool, bool, ptr, int, ptr, int),
get_data(bool, float, ptr, ptr, ptr), generate_frames(bool, bool, int, ptr), observations1/2, th.max, replicate + cat —
must give, bit for bit, what the fused stm_bfs gives and what the oracle gives.  The loop below restates the reference's
call sequence (the reference tree itself is not on the GPU box; tests/test_reference_main.py runs the reference's own file
against the same facade in the build container)."""
import os
import sys

import numpy as np
import pytest
import torch as th

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_style_bfs(AtariEnv, pool, step_env, state, tree_depth, gamma, A, n_frame_stack=4):
    """The call sequence of policies/cule_bfs.py:98-209 (max_width == -1), against a torchcule-style env object."""
    state_clone = state.clone()
    pool.states[0] = step_env.states[0]                                     # :110-112
    pool.ram[0] = step_env.ram[0]
    pool.frame_states[0] = step_env.frame_states[0]
    pool.rewards.zero_(); pool.observations1.zero_(); pool.observations2.zero_(); pool.done.zero_()     # :115-118
    actions0 = pool.action_set
    for depth in range(tree_depth):
        num_envs = A ** (depth + 1)
        pool.expand(num_envs)                                                # :143
        depth_actions = actions0.repeat(A ** depth)                          # :144
        for frame in range(pool.frameskip):
            super(AtariEnv, pool).step(pool.fire_reset and pool.is_training, False, depth_actions.data_ptr(), 0,
                                       pool.done.data_ptr(), 0)              # :154-155
            pool.get_data(pool.episodic_life, gamma ** depth, pool.done.data_ptr(), pool.rewards.data_ptr(),
                          pool.lives.data_ptr())                             # :157-158
            if frame == pool.frameskip - 2:
                pool.generate_frames(pool.rescale, False, pool.num_channels, pool.observations2[:num_envs].data_ptr())
            if frame == pool.frameskip - 1:
                pool.generate_frames(pool.rescale, False, pool.num_channels, pool.observations1[:num_envs].data_ptr())
        new_obs = th.max(pool.observations1[:num_envs], pool.observations2[:num_envs])      # :168
        new_obs = new_obs.squeeze(dim=-1).unsqueeze(dim=1)
        tmp = state_clone.reshape(state_clone.shape[0], -1).repeat(1, A).view(-1, 4 * 84 * 84)
        state_clone = tmp.reshape(-1, 4, 84, 84)                            # replicate_state :64-69
        state_clone = th.cat((state_clone[:, 1:n_frame_stack], new_obs), dim=1)             # :172
    rewards = pool.rewards[:num_envs].clone()
    pool.set_size(1)
    return state_clone, rewards, pool.done[:num_envs].clone(), pool.lives[:num_envs].clone(), pool.blobs[:num_envs].clone()


@pytest.mark.parametrize("game,depth", [("synth_breakout", 2), ("synth_breakout", 3), ("synth_pong", 2), ("synth_pacman", 2)])
def test_frame_by_frame_facade_equals_fused_bfs_and_oracle(game, depth):
    from oracle import orc
    from softtreemax_b200.atari import Env as AtariEnv
    g = orc.Game(game)
    A = g.n_actions
    roots, stacks = orc.make_roots(g, 3, seed=21, base_steps=70, spread=40)
    step_env = AtariEnv(num_envs=1, device="cuda:0", env_name=game)
    pool = AtariEnv(num_envs=A ** depth, device="cuda:0", env_name=game)
    pool.set_size(1)
    eng = pool.eng
    for r in range(3):
        step_env.blobs[:1] = th.from_numpy(roots[r:r + 1]).to(eng.device)
        root_stack = th.from_numpy(stacks[r:r + 1]).to(eng.device)
        leaves, rewards, done, lives, blobs = reference_style_bfs(AtariEnv, pool, step_env, root_stack, depth, 0.97, A)
        fused = eng.bfs(step_env.blobs[:1].contiguous(), root_stack, depth, gamma=0.97, clip_reward=False, want_states=True)
        th.cuda.synchronize()
        assert th.equal(leaves, fused["stacks"]), "leaf stacks: frame-by-frame facade vs stm_bfs"
        assert th.equal(rewards, fused["rewards"]) and th.equal(done, fused["done"]) and th.equal(lives, fused["lives"])
        assert th.equal(blobs, fused["states"]), "leaf emulator states"
        o_stack, o_rew, o_done, o_lives, o_st = g.bfs(roots[r:r + 1], stacks[r:r + 1], depth, gamma=0.97, clip_reward=False)
        assert np.array_equal(leaves.cpu().numpy(), o_stack[0]) and np.array_equal(rewards.cpu().numpy(), o_rew[0])
        assert np.array_equal(done.cpu().numpy(), o_done[0]) and np.array_equal(blobs.cpu().numpy(), o_st[0])


def test_frame_by_frame_freezes_like_the_fused_step():
    """An env that finishes inside an agent step still plays the rest of that step and is frozen from the next one on, in
    both forms (stm_step with frameskip 4 == four stm_step_frame calls with the done flags of the step's first frame)."""
    from oracle import orc
    from softtreemax_b200.atari import Env as AtariEnv
    from softtreemax_b200.engine import Engine
    g = orc.Game("synth_breakout")
    roots, _ = orc.make_roots(g, 64, seed=5, base_steps=200, spread=64)
    eng = Engine("synth_breakout")
    try:
        st = th.from_numpy(roots).to(eng.device)
        n = st.shape[0]
        rw, dn = th.zeros(n, device=eng.device), th.zeros(n, dtype=th.uint8, device=eng.device)
        lv = th.full((n,), 9, dtype=th.int32, device=eng.device)          # more lives than any root has: everyone "loses a life" in frame 0
        a = (th.arange(n, device=eng.device) % 4).to(th.uint8)
        s1, r1, d1, l1 = st.clone(), rw.clone(), dn.clone(), lv.clone()
        for _ in range(2):
            eng.step(s1, r1, d1, l1, actions=a, frameskip=4, weight=0.5, render=True)
        s2, r2, d2, l2 = st.clone(), rw.clone(), dn.clone(), lv.clone()
        codes = th.tensor(eng.action_codes, dtype=th.uint8, device=eng.device)[a.long()]
        for _ in range(2):
            frozen = d2.clone()
            for f in range(4):
                eng.step_frame(s2, codes, n, True, 0.5, r2, d2, l2, frozen)
        th.cuda.synchronize()
        assert d1.all() and th.equal(d1, d2)
        assert th.equal(s1, s2) and th.equal(r1, r2) and th.equal(l1, l2)
    finally:
        eng.close()
