"""stable_baselines3.common.vec_env.dummy_vec_env.DummyVecEnv (policies/cule_bfs.py:5 type-checks against it; PPO
wraps a single gym-style env in it): steps each env in turn, resets an env whose episode ended and keeps the final
observation in info["terminal_observation"]."""
import numpy as np

from .base_vec_env import VecEnv


def _scalar(x):
    """done / reward may arrive as Python scalars, numpy scalars or one-element tensors (CPU or GPU)."""
    return x.item() if hasattr(x, "item") else x


class DummyVecEnv(VecEnv):
    def __init__(self, env_fns):
        self.envs = [fn() for fn in env_fns]
        env = self.envs[0]
        super().__init__(len(self.envs), env.observation_space, env.action_space)
        shape = tuple(env.observation_space.shape)
        self.buf_obs = np.zeros((self.num_envs,) + shape, dtype=env.observation_space.dtype)
        self.buf_dones = np.zeros((self.num_envs,), dtype=bool)
        self.buf_rews = np.zeros((self.num_envs,), dtype=np.float32)
        self.buf_infos = [{} for _ in range(self.num_envs)]
        self.actions = None
        self.metadata = getattr(env, "metadata", {})

    def step_async(self, actions):
        self.actions = actions

    def step_wait(self):
        for i, env in enumerate(self.envs):
            obs, rew, done, info = env.step(self.actions[i])
            self.buf_rews[i], self.buf_dones[i], self.buf_infos[i] = float(_scalar(rew)), bool(_scalar(done)), info
            if self.buf_dones[i]:
                self.buf_infos[i]["terminal_observation"] = obs
                obs = env.reset()
            self.buf_obs[i] = obs
        return self.buf_obs.copy(), self.buf_rews.copy(), self.buf_dones.copy(), [dict(i) for i in self.buf_infos]

    def reset(self):
        for i, env in enumerate(self.envs):
            self.buf_obs[i] = env.reset()
        return self.buf_obs.copy()

    def close(self):
        for env in self.envs:
            try:
                env.close()
            except NotImplementedError:          # the reference's CuleEnv.close raises (cule_env.py:106-107)
                pass

    def get_attr(self, attr_name, indices=None):
        return [getattr(e, attr_name) for e in self.envs]

    def set_attr(self, attr_name, value, indices=None):
        for e in self.envs:
            setattr(e, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        return [getattr(e, method_name)(*args, **kwargs) for e in self.envs]
