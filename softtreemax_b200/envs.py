"""CuleEnv — mirror of environments/cule_env.py:14-107 (single root env with the DeepMind-style
wrappers), with the emulator on the GPU instead of CuLE's CPU back-end (cule_env.py:20).
No gym dependency (gym is not installable here): the class exposes the same attributes and the
reset/step contract SB3's DummyVecEnv expects."""
import math
from collections import deque

import numpy as np
import torch

from .atari import Env as AtariEnv, Rom as AtariRom, _Space


class CuleEnv:
    def __init__(self, device, env_kwargs, n_frame_stack=4, noop_max=30, clip_reward=True, fire_reset=True):
        self.device = device
        self.env_kwargs = env_kwargs
        actions = AtariRom(env_kwargs["env_name"]).minimal_actions()
        self.env = AtariEnv(num_envs=1, device="cuda:0" if str(device) in ("cpu", "auto") else device, **env_kwargs)
        self.env.reset(initial_steps=1, verbose=1)
        self.lives = 0
        self.life_termination = False
        self.n_frame_stack = n_frame_stack
        self.state_buffer = deque([], maxlen=n_frame_stack)
        self.training = True
        self.noop_max = noop_max
        self.clip_reward = clip_reward
        self.fire_reset = fire_reset
        self.reward_range = (-math.inf, math.inf)
        self.metadata = {"render.modes": ["human", "rgb_array"]}
        self.observation_space = _Space(shape=(n_frame_stack, 84, 84))
        self.action_space = _Space(n=len(actions))

    def _reset_buffer(self):
        for _ in range(self.n_frame_stack):
            self.state_buffer.append(torch.zeros(84, 84, dtype=torch.uint8, device=self.env.device))

    def _obs(self):
        return torch.stack(list(self.state_buffer), 0).cpu().numpy()                      # cule_env.py:72,93

    def obs_tensor(self):
        """The current observation stack (4, 84, 84) u8 ON THE GPU: what reset()/step() just returned as numpy, without the
        device->host copy.  A rollout that keeps its observations on the device reads this."""
        return torch.stack(list(self.state_buffer), 0)

    def reset(self):                                                                        # cule_env.py:43-54
        obs = self.reset_no_fire()
        if not self.fire_reset:
            return obs
        obs, _, done, _ = self.step(1)
        if done:
            self.reset_no_fire()
        obs, _, done, _ = self.step(2)
        if done:
            self.reset_no_fire()
        return obs

    def reset_no_fire(self):                                                                # cule_env.py:56-72
        obs = torch.zeros(84, 84, dtype=torch.uint8, device=self.env.device)          # cule_env.py:57
        if self.life_termination:
            self.life_termination = False
            self.env.step(torch.tensor([0]))      # a no-op after the lost life; its frame is NOT shown: a zero frame is (:58-60)
        else:
            self._reset_buffer()
            noops = np.random.randint(self.noop_max + 1)          # the globally seeded generator (utils.set_seed), :65
            obs = self.env.reset(initial_steps=noops, verbose=1)[0, :, :, 0]
            self.env.lives[0] = 5                                 # :67
        self.last_frame = obs
        self.state_buffer.append(obs)
        self.lives = int(self.env.lives[0].item())
        return self._obs()

    def step(self, action, want_obs=True):                                                  # cule_env.py:74-93
        """want_obs=False: the caller reads obs_tensor() instead; None is returned in place of the numpy stack."""
        obs, reward, done, info = self.env.step(torch.tensor([action]))
        info = {"ale.lives": info["ale.lives"], "orig_reward": float(reward[0].item())}
        if self.clip_reward:
            reward = torch.sign(reward)
        obs = obs[0, :, :, 0]
        self.state_buffer.append(obs)
        self.last_frame = obs
        lives = int(info["ale.lives"][0].item())
        done = bool(done[0].item())
        if self.training and lives < self.lives and lives > 0:       # lives > 0 for Q*bert (cule_env.py:88)
            self.life_termination = True
            done = True
        self.lives = lives
        return (self._obs() if want_obs else None), reward[0].cpu().numpy(), done, info

    def train(self):
        self.training = True

    def eval(self):
        self.training = False

    def render(self, mode="human"):
        raise NotImplementedError

    def close(self):
        pass


class CuleEnvMultiple:
    """Mirror of environments/cule_env_multiple.py:13-80 — the depth-0 arm: `n_envs` root environments
    stepped together (K2 + K3 with R roots and no tree), SB3 VecEnv conventions (batched numpy in/out,
    one info dict per env).  The stack of the last `n_frame_stack` observations lives on the GPU
    (`self.env.stacks`, shifted by K3) instead of in a deque of tensors (:27,40-42,49,69-74).

    An env whose episode ended is restarted on the next `step` from the power-on state + one no-op
    step, with a zeroed stack.  (The reference never resets an env here: it relies on torchcule's
    own restart of finished envs, which is not visible in the reference's sources.)
    """

    def __init__(self, device, env_kwargs, n_frame_stack=4, noop_max=30, clip_reward=True, fire_reset=True, n_envs=1):
        if n_frame_stack != 4:
            raise NotImplementedError("n_frame_stack = 4 (what main.py uses)")
        self.device, self.env_kwargs, self.num_envs = device, env_kwargs, n_envs
        actions = AtariRom(env_kwargs["env_name"]).minimal_actions()
        self.env = AtariEnv(num_envs=n_envs, device="cuda:0" if str(device) in ("cpu", "auto") else device, **env_kwargs)
        self.env.reset(initial_steps=1, verbose=1)
        self._fresh = (self.env.blobs.clone(), self.env.lives.clone())        # power-on + 1 no-op, for restarts
        self.lives = 0
        self.n_frame_stack, self.training = n_frame_stack, True
        self.noop_max, self.clip_reward, self.fire_reset = noop_max, clip_reward, fire_reset
        self.reward_range = (-math.inf, math.inf)
        self.metadata = {"render.modes": ["human", "rgb_array"]}
        self.observation_space = _Space(shape=(n_frame_stack, 84, 84))
        self.action_space = _Space(n=len(actions))
        self._ended = torch.zeros(n_envs, dtype=torch.bool, device=self.env.device)

    def _stack(self):
        return self.env.stacks.cpu().numpy()                                              # :50,74

    def reset(self):                                                                        # :44-50
        noops = np.random.randint(self.noop_max + 1)            # the globally seeded generator, cule_env_multiple.py:47
        self.env.reset(initial_steps=noops, verbose=1)
        self.env.lives[0] = 5                                   # :49 (env 0 only, as the reference writes it)
        self.env.stacks[:, :3] = 0                              # _reset_buffer(): three zero frames under the first observation
        self.last_frame = self.env.stacks[:, 3]
        self.lives = self.env.lives.cpu().numpy()
        self._ended.zero_()
        return self._stack()

    def step(self, action):                                                                 # :52-74
        env = self.env
        if bool(self._ended.any()):
            idx = self._ended.nonzero().reshape(-1)
            env.blobs[idx] = self._fresh[0][idx]
            env.lives[idx] = self._fresh[1][idx]
            env.stacks[idx] = 0
        obs, reward, done, info = env.step(torch.as_tensor(np.asarray(action)))
        self._ended = done.clone()
        lives = info["ale.lives"].cpu().numpy()
        orig, done_np = reward.cpu().numpy(), done.cpu().numpy()
        infos = [{"ale.lives": int(lives[i]), "orig_reward": orig[i], "done": done_np[i]} for i in range(self.num_envs)]
        if self.clip_reward:
            reward = torch.sign(reward)
        self.last_frame = obs[:, :, :, 0]
        self.lives = lives
        return self._stack(), reward.cpu().numpy(), done.float().cpu().numpy(), infos

    def train(self):
        self.training = True

    def eval(self):
        self.training = False

    def seed(self, seed=None):
        np.random.seed(seed)

    def render(self):
        raise NotImplementedError

    def close(self):
        pass
