"""stable_baselines3.PPO as the reference drives it (main.py:52-68,78,87-88): on-policy rollouts of n_steps per env with
policy.forward, GAE(lambda) in a rollout buffer, n_epochs passes over minibatches of batch_size with the clipped
surrogate + value + entropy loss through policy.evaluate_actions, Adam with the schedules of main.py:50-51,
gradient-norm clipping 0.5 — SB3 1.4.0's algorithm restated (OnPolicyAlgorithm.collect_rollouts, RolloutBuffer,
PPO.train), with the locals the reference's callback reads (callbacks.py:19-57: `infos`, `dones`, `self`)."""
import time

import numpy as np
import torch as th
from torch.nn import functional as F

from .common.callbacks import BaseCallback, CallbackList
from .common.logger import Logger
from .common.utils import explained_variance, get_device, obs_as_tensor, set_random_seed
from .common.vec_env import DummyVecEnv, VecEnv


def _schedule(v):
    return v if callable(v) else (lambda _: float(v))


class RolloutBuffer:
    def __init__(self, buffer_size, observation_space, n_envs, device, gamma, gae_lambda):
        self.buffer_size, self.n_envs, self.device, self.gamma, self.gae_lambda = buffer_size, n_envs, device, gamma, gae_lambda
        self.obs_shape = tuple(observation_space.shape)
        self.reset()

    def reset(self):
        n, e = self.buffer_size, self.n_envs
        self.observations = np.zeros((n, e) + self.obs_shape, dtype=np.float32)
        self.actions = np.zeros((n, e, 1), dtype=np.float32)
        self.rewards, self.returns, self.episode_starts, self.values, self.log_probs, self.advantages = (
            np.zeros((n, e), dtype=np.float32) for _ in range(6))
        self.pos, self.full = 0, False

    def add(self, obs, action, reward, episode_start, value, log_prob):
        self.observations[self.pos] = np.asarray(obs)
        self.actions[self.pos] = np.asarray(action).reshape(self.n_envs, 1)
        self.rewards[self.pos] = np.asarray(reward)
        self.episode_starts[self.pos] = np.asarray(episode_start)
        self.values[self.pos] = value.clone().cpu().numpy().flatten()
        self.log_probs[self.pos] = log_prob.clone().cpu().numpy().reshape(-1)
        self.pos += 1
        self.full = self.pos == self.buffer_size

    def compute_returns_and_advantage(self, last_values, dones):
        last_values = last_values.clone().cpu().numpy().flatten()
        last_gae = 0
        for step in reversed(range(self.buffer_size)):
            if step == self.buffer_size - 1:
                next_non_terminal, next_values = 1.0 - np.asarray(dones, dtype=np.float32), last_values
            else:
                next_non_terminal, next_values = 1.0 - self.episode_starts[step + 1], self.values[step + 1]
            delta = self.rewards[step] + self.gamma * next_values * next_non_terminal - self.values[step]
            last_gae = delta + self.gamma * self.gae_lambda * next_non_terminal * last_gae
            self.advantages[step] = last_gae
        self.returns = self.advantages + self.values

    def get(self, batch_size):
        n = self.buffer_size * self.n_envs
        flat = lambda a: a.swapaxes(0, 1).reshape(n, *a.shape[2:])
        data = [flat(a) for a in (self.observations, self.actions, self.values[..., None], self.log_probs[..., None],
                                  self.advantages[..., None], self.returns[..., None])]
        idx = np.random.permutation(n)
        for s in range(0, n, batch_size):
            b = idx[s:s + batch_size]
            o, a, v, lp, adv, ret = (th.as_tensor(d[b]).to(self.device) for d in data)
            yield o, a, v.flatten(), lp.flatten(), adv.flatten(), ret.flatten()


class PPO:
    def __init__(self, policy, env, learning_rate=3e-4, n_steps=2048, batch_size=64, n_epochs=10, gamma=0.99, gae_lambda=0.95,
                 clip_range=0.2, clip_range_vf=None, ent_coef=0.0, vf_coef=0.5, max_grad_norm=0.5, use_sde=False,
                 target_kl=None, tensorboard_log=None, create_eval_env=False, policy_kwargs=None, verbose=0, seed=None,
                 device="auto", _init_setup_model=True):
        self.device = get_device(device)
        self.verbose, self.seed = verbose, seed
        self.env = env if isinstance(env, VecEnv) else DummyVecEnv([lambda: env])
        self.n_envs = self.env.num_envs
        self.observation_space, self.action_space = self.env.observation_space, self.env.action_space
        self.policy_class, self.policy_kwargs = policy, dict(policy_kwargs or {})
        self.learning_rate, self.lr_schedule = learning_rate, _schedule(learning_rate)
        self.clip_range = _schedule(clip_range)
        self.clip_range_vf = None if clip_range_vf is None else _schedule(clip_range_vf)
        self.n_steps, self.batch_size, self.n_epochs = n_steps, batch_size, n_epochs
        self.gamma, self.gae_lambda = gamma, gae_lambda
        self.ent_coef, self.vf_coef, self.max_grad_norm, self.use_sde, self.target_kl = ent_coef, vf_coef, max_grad_norm, use_sde, target_kl
        self.num_timesteps, self._total_timesteps, self._n_updates = 0, 0, 0
        self._current_progress_remaining = 1.0
        self._last_obs, self._last_episode_starts = None, None
        self.logger = Logger(verbose)
        self.start_time = None
        if _init_setup_model:
            self._setup_model()

    def _setup_model(self):
        if self.seed is not None:
            set_random_seed(self.seed)
        self.rollout_buffer = RolloutBuffer(self.n_steps, self.observation_space, self.n_envs, self.device, self.gamma, self.gae_lambda)
        self.policy = self.policy_class(self.observation_space, self.action_space, self.lr_schedule, use_sde=self.use_sde,
                                        **self.policy_kwargs)
        self.policy = self.policy.to(self.device)

    def get_env(self):
        return self.env

    def predict(self, observation, state=None, mask=None, deterministic=False):
        return self.policy.predict(observation, state, mask, deterministic)

    # ------------------------------------------------------------------------------------------------------------------
    def collect_rollouts(self, env, callback, rollout_buffer, n_rollout_steps):
        self.policy.set_training_mode(False)
        n_steps = 0
        rollout_buffer.reset()
        callback.on_rollout_start()
        while n_steps < n_rollout_steps:
            with th.no_grad():
                obs_tensor = obs_as_tensor(self._last_obs, self.device)
                actions, values, log_probs = self.policy.forward(obs_tensor)
            actions = actions.cpu().numpy()
            clipped_actions = actions
            new_obs, rewards, dones, infos = env.step(clipped_actions)
            self.num_timesteps += env.num_envs
            callback.update_locals(locals())
            if callback.on_step() is False:
                return False
            n_steps += 1
            actions = actions.reshape(-1, 1)
            rollout_buffer.add(self._last_obs, actions, rewards, self._last_episode_starts, values, log_probs)
            self._last_obs, self._last_episode_starts = new_obs, dones
        with th.no_grad():
            values = self.policy.predict_values(obs_as_tensor(new_obs, self.device))
        rollout_buffer.compute_returns_and_advantage(last_values=values, dones=dones)
        callback.on_rollout_end()
        return True

    def train(self):
        self.policy.set_training_mode(True)
        lr = self.lr_schedule(self._current_progress_remaining)
        self.logger.record("train/learning_rate", lr)
        for g in self.policy.optimizer.param_groups:
            g["lr"] = lr
        clip_range = self.clip_range(self._current_progress_remaining)
        entropy_losses, pg_losses, value_losses, clip_fractions = [], [], [], []
        for _ in range(self.n_epochs):
            for observations, actions, old_values, old_log_prob, advantages, returns in self.rollout_buffer.get(self.batch_size):
                actions = actions.long().flatten()
                values, log_prob, entropy = self.policy.evaluate_actions(observations, actions)
                values = values.flatten()
                advantages = (advantages - advantages.mean()) / (advantages.std() + 1e-8)
                ratio = th.exp(log_prob - old_log_prob)
                policy_loss = -th.min(advantages * ratio, advantages * th.clamp(ratio, 1 - clip_range, 1 + clip_range)).mean()
                pg_losses.append(policy_loss.item())
                clip_fractions.append(th.mean((th.abs(ratio - 1) > clip_range).float()).item())
                values_pred = values
                if self.clip_range_vf is not None:
                    cv = self.clip_range_vf(self._current_progress_remaining)
                    values_pred = old_values + th.clamp(values - old_values, -cv, cv)
                value_loss = F.mse_loss(returns, values_pred)
                value_losses.append(value_loss.item())
                entropy_loss = -th.mean(entropy) if entropy is not None else -th.mean(-log_prob)
                entropy_losses.append(entropy_loss.item())
                loss = policy_loss + self.ent_coef * entropy_loss + self.vf_coef * value_loss
                self.policy.optimizer.zero_grad()
                loss.backward()
                th.nn.utils.clip_grad_norm_(self.policy.parameters(), self.max_grad_norm)
                self.policy.optimizer.step()
        self._n_updates += self.n_epochs
        ev = explained_variance(self.rollout_buffer.values.flatten(), self.rollout_buffer.returns.flatten())
        self.logger.record("train/entropy_loss", float(np.mean(entropy_losses)))
        self.logger.record("train/policy_gradient_loss", float(np.mean(pg_losses)))
        self.logger.record("train/value_loss", float(np.mean(value_losses)))
        self.logger.record("train/clip_fraction", float(np.mean(clip_fractions)))
        self.logger.record("train/loss", float(loss.item()))
        self.logger.record("train/explained_variance", float(ev))
        self.logger.record("train/n_updates", self._n_updates)
        self.logger.record("train/clip_range", clip_range)

    def learn(self, total_timesteps, callback=None, log_interval=1, eval_env=None, eval_freq=-1, n_eval_episodes=5,
              tb_log_name="PPO", eval_log_path=None, reset_num_timesteps=True):
        self.start_time = time.time()
        if reset_num_timesteps:
            self.num_timesteps = 0
        self._total_timesteps = total_timesteps
        if self._last_obs is None or reset_num_timesteps:
            self._last_obs = self.env.reset()
            self._last_episode_starts = np.ones((self.env.num_envs,), dtype=bool)
        if isinstance(callback, list):
            callback = CallbackList(callback)
        elif callback is None:
            callback = BaseCallback()
        callback.init_callback(self)
        iteration = 0
        callback.on_training_start(locals(), globals())
        while self.num_timesteps < total_timesteps:
            if not self.collect_rollouts(self.env, callback, self.rollout_buffer, n_rollout_steps=self.n_steps):
                break
            iteration += 1
            self._current_progress_remaining = 1.0 - float(self.num_timesteps) / float(total_timesteps)
            if log_interval is not None and iteration % log_interval == 0:
                fps = int(self.num_timesteps / max(time.time() - self.start_time, 1e-9))
                self.logger.record("time/iterations", iteration)
                self.logger.record("time/fps", fps)
                self.logger.record("time/total_timesteps", self.num_timesteps)
                self.logger.dump(step=self.num_timesteps)
            self.train()
        callback.on_training_end()
        return self
