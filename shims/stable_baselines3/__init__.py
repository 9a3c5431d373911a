"""Import-compatible stand-in for the parts of Stable-Baselines3 1.4.0 that the reference uses (main.py:11-13,
policies/actor_critic_ts.py:7, policies/actor_critic_depth0.py:6, policies/cule_bfs.py:5, callbacks.py:8,
environments/cule_env_multiple.py:10).  SB3 (pinned `stable-baselines3==1.4.0`, reference Dockerfile:10) is a
third-party dependency that cannot be installed here; its published algorithm — PPO with GAE, clipped surrogate,
Adam(eps=1e-5), gradient-norm clipping 0.5 — and the class surface the reference touches are restated:

    stable_baselines3.PPO
    stable_baselines3.common.utils            get_device, get_linear_fn, obs_as_tensor, set_random_seed
    stable_baselines3.common.evaluation       evaluate_policy
    stable_baselines3.common.policies         ActorCriticCnnPolicy (NatureCNN, mlp_extractor, action_dist, ...)
    stable_baselines3.common.callbacks        BaseCallback, CallbackList
    stable_baselines3.common.vec_env          VecEnv, DummyVecEnv
    stable_baselines3.common.logger           Logger (name_to_value, record, dump)

With `shims/` on the path the reference's own main.py, policies/*.py, environments/*.py and callbacks.py run
unmodified."""
from .ppo import PPO  # noqa: F401

__version__ = "1.4.0+stm_b200_shim"
