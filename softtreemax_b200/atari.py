"""Host-side mirror of the `torchcule.atari` surface that the reference touches
(policies/cule_bfs.py:3-4,29-31,55-62; environments/cule_env.py:7-8,18-22,60-71,76-93), backed by
libstm_b200.so.  Everything lives on the GPU: there is no CPU back-end and no CPU->GPU crossover
(the reference's CROSSOVER_DICT / copy_envs, cule_bfs.py:7,71-79,133-137, has nothing to do here).
"""
import numpy as np
import torch

from . import games
from .engine import Engine

_ENGINES = {}


def engine_for(env_name, device=0):
    """One Engine (one stm_ctx) per (cartridge, device), shared by every Env / CuleBFS on it."""
    key = (games.resolve(env_name), int(device))
    if key not in _ENGINES:
        _ENGINES[key] = Engine(key[0], device=key[1])
    return _ENGINES[key]


class Rom:
    """torchcule.atari.Rom: `Rom(env_name).minimal_actions()` (cule_bfs.py:29-30)."""

    def __init__(self, env_name):
        self.desc, self.rom, self.info = games.load(env_name)

    def minimal_actions(self):
        return torch.tensor(self.info["actions"], dtype=torch.uint8)


class _Space:
    def __init__(self, n=None, shape=None):
        self.n, self.shape = n, shape


class Env:
    """torchcule.atari.Env for the attributes and methods the reference uses.

    Tensors (all on the GPU): `blobs` (N, 256) u8 is the storage; `ram` = blobs[:, :128] and
    `states` = blobs[:, 128:] are views of it (`frame_states` is zero-width: the TIA needs no separate
    render state), so the reference's root injection `pool.states[0] = env.states[0]` etc.
    (cule_bfs.py:110-112) keeps working as plain tensor assignment.
    """

    def __init__(self, num_envs=1, device="cuda:0", action_set=None, env_name="BreakoutNoFrameskip-v4",
                 color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4, **_):
        if color_mode != "gray" or not rescale or repeat_prob != 0.0:
            raise NotImplementedError("only color_mode='gray', rescale=True, repeat_prob=0.0 (what main.py:37-38 uses)")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("softtreemax_b200 has no CPU back-end: the emulator runs on the GPU only")
        self.eng = engine_for(env_name, dev.index or 0)
        self.device, self.is_cuda = self.eng.device, True
        self.env_name, self.num_envs = env_name, num_envs
        self.episodic_life, self.frameskip, self.rescale, self.num_channels = episodic_life, frameskip, rescale, 1
        self.fire_reset, self.is_training = False, True
        self.action_set = torch.tensor(self.eng.action_codes, dtype=torch.uint8, device=self.device)
        self.action_space = _Space(n=self.eng.n_actions)
        self.observation_space = _Space(shape=(1, 84, 84))
        self.blobs, self.rewards, self.done, self.lives = self.eng.reset_states(num_envs)
        self.stacks = torch.zeros((num_envs, 4, 84, 84), dtype=torch.uint8, device=self.device)
        self._n = num_envs

    # views the reference indexes directly
    @property
    def ram(self):
        return self.blobs[:, :128]

    @property
    def states(self):
        return self.blobs[:, 128:]

    @property
    def frame_states(self):
        return self.blobs[:, 256:]

    def size(self):
        return self._n

    def set_size(self, n):
        assert n <= self.num_envs
        self._n = n

    def sync_other_stream(self):      # everything runs on the caller's stream: nothing to fence
        pass

    sync_this_stream = update_frame_states = sync_other_stream

    def reset(self, initial_steps=0, verbose=False, **_):
        """Power-on state, then `initial_steps` NOOP agent steps; returns obs (N, 84, 84, 1) u8."""
        self.blobs, self.rewards, self.done, self.lives = self.eng.reset_states(self.num_envs)
        self.stacks.zero_()
        self._started = False
        noop = torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device)
        for _ in range(max(1, int(initial_steps))):
            self._raw_step(noop)
        return self.stacks[:, 3].unsqueeze(-1)

    def _raw_step(self, actions):
        self.done.zero_()
        self.rewards.zero_()
        self.eng.step(self.blobs, self.rewards, self.done, self.lives, actions=actions, frameskip=self.frameskip,
                      episodic_life=self.episodic_life, weight=1.0, render=True)
        self.stacks = self.eng.frames(self.stacks, self.num_envs, fanout=1)

    def step(self, actions):
        """actions: policy-action indices.  -> (obs (N,84,84,1) u8, reward (N,) f32, done (N,) bool, info)"""
        a = torch.as_tensor(actions).to(self.device, torch.uint8).reshape(-1)
        self._raw_step(a)
        return (self.stacks[:, 3].unsqueeze(-1), self.rewards.clone(), self.done.bool(),
                {"ale.lives": self.lives.clone()})
