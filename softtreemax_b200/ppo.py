"""Minimal PPO driver for the SoftTreeMax policy — the caller on the other side of the hot path
(SURVEY §8f rows 1-2).  The reference uses Stable-Baselines3 1.4.0's PPO (main.py:52-68,78), a
third-party dependency that is not installable here; this restates its algorithm for the single-env
case the tree path uses (n_envs = 1, main.py:47): rollout of n_steps with policy.forward, GAE(lambda),
n_epochs x minibatches of clipped-surrogate + value + entropy loss through policy.evaluate_actions,
Adam(eps=1e-5) with linearly decayed learning rate and clip range, gradient-norm clipping 0.5.
"""
import numpy as np
import torch as th

from . import shard


def get_linear_fn(start, end, end_fraction):
    """SB3's schedule: value as a function of progress_remaining (1 -> 0)."""
    def fn(progress_remaining):
        if (1 - progress_remaining) > end_fraction:
            return end
        return start + (1 - progress_remaining) * (end - start) / end_fraction
    return fn


class PPO:
    def __init__(self, policy, env, learning_rate=2.5e-4, n_steps=128, batch_size=32, n_epochs=3, gamma=0.99,
                 gae_lambda=0.95, clip_range=0.1, ent_coef=0.01, vf_coef=1.0, max_grad_norm=0.5,
                 policy_kwargs=None, verbose=0, seed=None):
        self.env = env
        self.lr = learning_rate if callable(learning_rate) else (lambda _: learning_rate)
        self.clip = clip_range if callable(clip_range) else (lambda _: clip_range)
        self.n_steps, self.batch_size, self.n_epochs = n_steps, batch_size, n_epochs
        self.gamma, self.gae_lambda = gamma, gae_lambda
        self.ent_coef, self.vf_coef, self.max_grad_norm = ent_coef, vf_coef, max_grad_norm
        self.verbose = verbose
        self.rng = np.random.RandomState(seed)
        self.policy = policy(env.observation_space, env.action_space, self.lr, **(policy_kwargs or {}))
        self.num_timesteps = 0
        self.log = []
        self._step_cb = None

    def _rollout(self, obs, episode_start):
        """n_steps of one root env.  When the env can hand its observation over on the GPU (CuleEnv.obs_tensor) the rollout
        never moves an observation to the host: obs, the rollout buffer's observations and the policy input stay device
        tensors (the reference copies every observation device -> host -> device, cule_env.py:93 + SB3's obs_as_tensor)."""
        n = self.n_steps
        on_dev = hasattr(self.env, "obs_tensor")
        dev = self.policy.device
        buf = dict(obs=(th.zeros((n, 4, 84, 84), dtype=th.uint8, device=dev) if on_dev else np.zeros((n, 4, 84, 84), np.uint8)),
                   act=np.zeros(n, np.int64), rew=np.zeros(n, np.float32),
                   start=np.zeros(n, np.float32), val=th.zeros(n, device=dev), logp=th.zeros(n, device=dev))
        ep_rewards = []
        cur = self.env.obs_tensor() if on_dev else obs
        for t in range(n):
            with th.no_grad():
                a, v, lp = self.policy.forward(cur[None] if on_dev else th.from_numpy(cur).float()[None])
            buf["obs"][t] = cur
            buf["val"][t], buf["logp"][t] = v.reshape(-1)[0], lp.reshape(-1)[0]          # device-side: no sync
            a = int(a)                                                                   # the one sync of the step: the env needs the action
            buf["act"][t], buf["start"][t] = a, float(episode_start)
            if on_dev:
                _, r, done, info = self.env.step(a, want_obs=False)
            else:
                cur, r, done, info = self.env.step(a)
            if self._step_cb is not None:
                self._step_cb(self, [info], [done])
            buf["rew"][t] = float(r)
            self._ep_ret += float(info.get("orig_reward", r))
            self.num_timesteps += 1
            episode_start = done
            if done:
                ep_rewards.append(self._ep_ret)
                self._ep_ret = 0.0
                cur = self.env.reset()
            if on_dev:
                cur = self.env.obs_tensor()
        obs = cur
        with th.no_grad():
            o1 = obs[None] if on_dev else th.from_numpy(obs)[None]
            if getattr(self.policy, "use_leaves_v", False):
                # the stored val[t] are tree values (max over leaves): bootstrap with the same definition.  The tree of
                # this observation stays in the leaf cache and serves the first step of the next rollout.
                last_val = float(self.policy.forward(o1.float())[1].reshape(-1)[0])
            else:                                            # SB3: predict_values = the depth-0 value of the last obs
                feats = self.policy.extract_features(o1.to(dev))
                last_val = float(self.policy.value_net(feats).reshape(-1)[0])
        buf["val"], buf["logp"] = buf["val"].cpu().numpy(), buf["logp"].cpu().numpy()
        adv = np.zeros(n, np.float32)
        gae = 0.0
        for t in reversed(range(n)):                         # GAE(lambda), SB3 RolloutBuffer.compute_returns_and_advantage
            nonterminal = 1.0 - (float(episode_start) if t == n - 1 else buf["start"][t + 1])
            next_val = last_val if t == n - 1 else buf["val"][t + 1]
            delta = buf["rew"][t] + self.gamma * next_val * nonterminal - buf["val"][t]
            gae = delta + self.gamma * self.gae_lambda * nonterminal * gae
            adv[t] = gae
        buf["adv"], buf["ret"] = adv, adv + buf["val"]
        if on_dev:
            obs = obs.cpu().numpy()                          # the caller (learn) carries the last observation as numpy
        return buf, obs, episode_start, ep_rewards

    def _rollout_vec(self, obs, episode_start):
        """Rollout over a vectorised env (`env.num_envs` envs that restart themselves: CuleEnvMultiple, the depth-0
        arm).  Same buffer keys as _rollout, flattened to (n_steps * n_envs, ...)."""
        n, N = self.n_steps, self.env.num_envs
        buf = dict(obs=np.zeros((n, N, 4, 84, 84), np.uint8), act=np.zeros((n, N), np.int64), rew=np.zeros((n, N), np.float32),
                   start=np.zeros((n, N), np.float32), val=np.zeros((n, N), np.float32), logp=np.zeros((n, N), np.float32))
        ep_rewards = []
        for t in range(n):
            with th.no_grad():
                a, v, lp = self.policy.forward(th.from_numpy(obs))
            a = a.cpu().numpy()
            buf["obs"][t], buf["act"][t], buf["start"][t] = obs, a, episode_start
            buf["val"][t], buf["logp"][t] = v.reshape(-1).cpu().numpy(), lp.reshape(-1).cpu().numpy()
            obs, r, done, infos = self.env.step(a)
            if self._step_cb is not None:
                self._step_cb(self, infos, done)
            buf["rew"][t] = r
            self._ep_ret = self._ep_ret + np.array([float(i.get("orig_reward", x)) for i, x in zip(infos, r)], np.float64)
            self.num_timesteps += N
            episode_start = np.asarray(done, np.float32)
            for i in np.nonzero(episode_start)[0]:
                ep_rewards.append(float(self._ep_ret[i]))
                self._ep_ret[i] = 0.0
        with th.no_grad():
            last_val = self.policy.forward(th.from_numpy(obs))[1].reshape(-1).cpu().numpy()
        adv = np.zeros((n, N), np.float32)
        gae = np.zeros(N, np.float32)
        for t in reversed(range(n)):
            nonterminal = 1.0 - (episode_start if t == n - 1 else buf["start"][t + 1])
            next_val = last_val if t == n - 1 else buf["val"][t + 1]
            delta = buf["rew"][t] + self.gamma * next_val * nonterminal - buf["val"][t]
            gae = delta + self.gamma * self.gae_lambda * nonterminal * gae
            adv[t] = gae
        buf["adv"], buf["ret"] = adv, adv + buf["val"]
        buf = {k: v.reshape((n * N,) + v.shape[2:]) for k, v in buf.items()}
        return buf, obs, episode_start, ep_rewards

    def _train(self, buf, progress_remaining):
        pol = self.policy
        lr, clip = self.lr(progress_remaining), self.clip(progress_remaining)
        for g in pol.optimizer.param_groups:
            g["lr"] = lr
        dev = pol.device
        stats = []
        for _ in range(self.n_epochs):
            total = len(buf["act"])
            perm = self.rng.permutation(total)
            for s in range(0, total, self.batch_size):
                idx = perm[s:s + self.batch_size]
                obs = buf["obs"][th.from_numpy(idx).to(buf["obs"].device)] if isinstance(buf["obs"], th.Tensor) else th.from_numpy(buf["obs"][idx]).float()
                act = th.from_numpy(buf["act"][idx]).to(dev)
                adv = th.from_numpy(buf["adv"][idx]).to(dev)
                adv = (adv - adv.mean()) / (adv.std() + 1e-8)
                values, logp, entropy = pol.evaluate_actions(obs, act)
                ratio = th.exp(logp - th.from_numpy(buf["logp"][idx]).to(dev))
                pg = -th.min(adv * ratio, adv * th.clamp(ratio, 1 - clip, 1 + clip)).mean()
                vl = th.nn.functional.mse_loss(th.from_numpy(buf["ret"][idx]).to(dev), values.reshape(-1))
                ent = -entropy.mean()
                loss = pg + self.ent_coef * ent + self.vf_coef * vl
                pol.optimizer.zero_grad()
                loss.backward()
                shard.allreduce_gradients(pol.parameters())      # data-parallel ranks (no-op on one process)
                th.nn.utils.clip_grad_norm_(pol.parameters(), self.max_grad_norm)
                pol.optimizer.step()
                stats.append((float(loss.detach()), float(pg.detach()), float(vl.detach()), float(-ent.detach())))
        return np.mean(stats, axis=0)

    def learn(self, total_timesteps, callback=None):
        vec = hasattr(self.env, "num_envs")
        self._step_cb = getattr(callback, "on_step", None)       # per-env-step hook (callbacks.TrainingCallback)
        obs = self.env.reset()
        if vec:
            episode_start, self._ep_ret = np.ones(self.env.num_envs, np.float32), np.zeros(self.env.num_envs, np.float64)
        else:
            episode_start, self._ep_ret = True, 0.0
        while self.num_timesteps < total_timesteps:
            buf, obs, episode_start, ep = (self._rollout_vec if vec else self._rollout)(obs, episode_start)
            st = self._train(buf, 1.0 - self.num_timesteps / total_timesteps)
            rec = dict(timesteps=self.num_timesteps, loss=st[0], pg_loss=st[1], value_loss=st[2], entropy=st[3],
                       episodes=len(ep), ep_reward_mean=float(np.mean(ep)) if ep else None,
                       beta=float(getattr(self.policy, "beta", 1.0)), alpha=float(getattr(self.policy, "alpha", 1.0)))
            self.log.append(rec)
            if self.verbose:
                print(rec, flush=True)
            if callback is not None:
                callback(self, rec)
        return self


def evaluate_policy(model, env, n_eval_episodes=10, deterministic=False, return_episode_rewards=False):
    """stable_baselines3.common.evaluation.evaluate_policy for the two env kinds of this package (main.py:94-95):
    runs `n_eval_episodes` episodes with model.policy.predict and returns the un-clipped episode scores and
    lengths (or their mean / std)."""
    policy = model.policy if hasattr(model, "policy") else model
    scores, lengths = [], []
    if hasattr(env, "num_envs"):
        N = env.num_envs
        obs, ret, ln = env.reset(), np.zeros(N), np.zeros(N, np.int64)
        while len(scores) < n_eval_episodes:
            a, _ = policy.predict(obs, deterministic=deterministic)
            obs, r, done, infos = env.step(a)
            ret += np.array([float(i.get("orig_reward", x)) for i, x in zip(infos, r)])
            ln += 1
            for i in np.nonzero(done)[0]:
                scores.append(float(ret[i]))
                lengths.append(int(ln[i]))
                ret[i], ln[i] = 0.0, 0
    else:
        for _ in range(n_eval_episodes):
            obs, done, ret, ln = env.reset(), False, 0.0, 0
            while not done:
                a, _ = policy.predict(obs[None], deterministic=deterministic)
                obs, r, done, info = env.step(int(np.asarray(a).reshape(-1)[0]))
                ret += float(info.get("orig_reward", r))
                ln += 1
            scores.append(ret)
            lengths.append(ln)
    scores, lengths = scores[:n_eval_episodes], lengths[:n_eval_episodes]
    if return_episode_rewards:
        return scores, lengths
    return float(np.mean(scores)), float(np.std(scores))
