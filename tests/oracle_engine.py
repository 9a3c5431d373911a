"""TEST DOUBLE.  An object with the Engine methods the torchcule facade (softtreemax_b200/atari.py) calls, backed by the
CPU oracle (oracle/stm_oracle.c) on CPU tensors.  It exists so that the GPU-less container can run the REFERENCE's own
main.py / policies / environments on top of the shims (tests/test_reference_main.py) and check the plumbing of the
facade — pointer conventions, in-place expand, frame-by-frame stepping — against the oracle's whole-tree bfs.  Only tests
install it (atari._ENGINE_FACTORY); the product package never sees it and has no CPU path."""
import ctypes as C

import numpy as np
import torch

from oracle import orc


def _np(t):
    return t.numpy() if isinstance(t, torch.Tensor) else t


def _from_ptr(ptr, ctype, n):
    return np.ctypeslib.as_array((ctype * n).from_address(int(ptr)))


class OracleEngine:
    def __init__(self, game, device=0, scratch_envs=None):
        from softtreemax_b200 import games
        self.desc, self.rom, self.info = games.load(game)
        self.game = orc.Game(game)
        self.device = torch.device("cpu")
        self.n_actions = self.game.n_actions
        self.action_codes = list(self.game.desc.actions[:self.n_actions])
        self.max_step_envs = 1 << 30
        self._fb = {}                      # env index -> newest native frame (set by step / step_frame)
        self._obs = None

    def set_scratch_envs(self, n):
        pass

    def reset_states(self, n):
        s = np.stack([self.game.reset() for _ in range(n)]) if n else np.zeros((0, 256), np.uint8)
        return (torch.from_numpy(s), torch.zeros(n), torch.zeros(n, dtype=torch.uint8), torch.zeros(n, dtype=torch.int32))

    def expand(self, states, rewards, done, lives, fanout):
        rep = lambda t: t.repeat_interleave(fanout, dim=0).contiguous()
        return rep(states), rep(rewards), rep(done), rep(lives)

    def step(self, states, rewards, done, lives, actions=None, frameskip=4, episodic_life=True, weight=1.0, render=True):
        n = states.shape[0]
        s, r, d, l = _np(states), _np(rewards), _np(done), _np(lives)
        self._obs = np.zeros((n, 84, 84), np.uint8)
        scratch = np.zeros(2 * orc.NATIVE_BYTES, np.uint8)
        L = orc.lib()
        for i in range(n):
            a = int(actions[i]) if actions is not None else i % self.n_actions
            L.orc_agent_step(self.game.h, orc._p(s[i]), a, frameskip, int(episodic_life), C.c_float(weight),
                             orc._p(d[i:i + 1]), orc._p(r[i:i + 1]), orc._p(l[i:i + 1]), orc._p(self._obs[i]), orc._p(scratch))

    def frames(self, parent_stacks, n, fanout=1, out=None):
        if out is None:
            out = torch.empty((n, 4, 84, 84), dtype=torch.uint8)
        par = parent_stacks[:(n + fanout - 1) // fanout].repeat_interleave(fanout, dim=0)[:n]
        new = torch.cat([par[:, 1:], torch.from_numpy(self._obs[:n])[:, None]], dim=1)
        out.copy_(new)
        return out

    def step_frame(self, states, action_codes, n, episodic_life, weight, rewards, done, lives, frozen=None):
        s = _np(states)
        codes = _from_ptr(action_codes, C.c_uint8, n) if isinstance(action_codes, int) else _np(action_codes)
        r = _from_ptr(rewards, C.c_float, n) if isinstance(rewards, int) else _np(rewards)
        d = _from_ptr(done, C.c_uint8, n) if isinstance(done, int) else _np(done)
        l = _from_ptr(lives, C.c_int32, n) if isinstance(lives, int) else _np(lives)
        fz = _np(frozen) if frozen is not None else d.copy()
        self._native = np.zeros((n, orc.NATIVE_BYTES), np.uint8)
        self._stepped = np.zeros(n, bool)
        L = orc.lib()
        for i in range(n):
            if fz[i]:
                continue
            self._stepped[i] = True
            L.orc_frame(self.game.h, orc._p(s[i]), int(codes[i]), orc._p(self._native[i]))
            L.orc_get_data(self.game.h, orc._p(s[i]), int(episodic_life), C.c_float(weight), orc._p(d[i:i + 1]),
                           orc._p(r[i:i + 1]), orc._p(l[i:i + 1]))

    def frame_obs(self, obs, n):
        o = _from_ptr(obs, C.c_uint8, n * 84 * 84).reshape(n, 84, 84) if isinstance(obs, int) else _np(obs).reshape(n, 84, 84)
        for i in range(n):
            o[i] = orc.render_obs(self._native[i]) if self._stepped[i] else 0

    def bfs(self, root_states, root_stacks, depth, gamma=0.99, frameskip=4, episodic_life=True, clip_reward=True,
            want_states=False, out=None):
        stack, rew, done, lives, st = self.game.bfs(_np(root_states), _np(root_stacks), depth, gamma, frameskip, episodic_life,
                                                    clip_reward, want_state=want_states)
        t = torch.from_numpy
        return dict(stacks=t(stack.reshape(-1, 4, 84, 84)), rewards=t(rew.reshape(-1)), done=t(done.reshape(-1)),
                    lives=t(lives.reshape(-1)), states=t(st.reshape(-1, 256)) if want_states else None)
