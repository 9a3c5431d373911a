"""stable_baselines3.common.vec_env.VecEnv: the base class environments/cule_env_multiple.py:13 derives from."""


class VecEnv:
    def __init__(self, num_envs=1, observation_space=None, action_space=None):
        self.num_envs, self.observation_space, self.action_space = num_envs, observation_space, action_space

    def reset(self):
        raise NotImplementedError

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        return self.step(self._actions)

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        pass

    def seed(self, seed=None):
        return [None] * self.num_envs

    def get_attr(self, attr_name, indices=None):
        return [getattr(self, attr_name)] * self.num_envs

    def set_attr(self, attr_name, value, indices=None):
        setattr(self, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        return [getattr(self, method_name)(*args, **kwargs)]

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * self.num_envs

    @property
    def unwrapped(self):
        return self
