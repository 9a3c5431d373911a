"""stable_baselines3.common.evaluation.evaluate_policy (main.py:13,93-95): runs n_eval_episodes episodes through
model.predict on a (vectorised) env and returns the episode rewards and lengths, or their mean and std."""
import numpy as np

from .vec_env import DummyVecEnv, VecEnv


def evaluate_policy(model, env, n_eval_episodes=10, deterministic=True, render=False, callback=None, reward_threshold=None,
                    return_episode_rewards=False, warn=True):
    if not isinstance(env, VecEnv):
        env = DummyVecEnv([lambda: env])
    n_envs = env.num_envs
    episode_rewards, episode_lengths = [], []
    counts = np.zeros(n_envs, dtype=int)
    targets = np.array([(n_eval_episodes + i) // n_envs for i in range(n_envs)], dtype=int)
    cur_rew, cur_len = np.zeros(n_envs), np.zeros(n_envs, dtype=int)
    obs, states = env.reset(), None
    while (counts < targets).any():
        actions, states = model.predict(obs, state=states, deterministic=deterministic)
        if hasattr(actions, "cpu"):
            actions = actions.cpu().numpy()          # the reference's tree policy returns a tensor (actor_critic_ts.py:272-273)
        obs, rewards, dones, infos = env.step(np.asarray(actions).reshape(n_envs, *np.asarray(actions).shape[1:]))
        cur_rew += np.asarray(rewards, dtype=float).reshape(n_envs)
        cur_len += 1
        for i in range(n_envs):
            if counts[i] < targets[i] and bool(np.asarray(dones).reshape(n_envs)[i]):
                episode_rewards.append(float(cur_rew[i]))
                episode_lengths.append(int(cur_len[i]))
                counts[i] += 1
                cur_rew[i], cur_len[i] = 0.0, 0
                if callback is not None:
                    callback(locals(), globals())
    if return_episode_rewards:
        return episode_rewards, episode_lengths
    return float(np.mean(episode_rewards)), float(np.std(episode_rewards))
