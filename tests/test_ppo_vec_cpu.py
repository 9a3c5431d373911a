"""Host logic of the depth-0 arm without a GPU: the vectorised PPO rollout / GAE /
minibatch update over a stand-in VecEnv, with the depth-0 CNN policy on the CPU."""
import numpy as np
import torch

from softtreemax_b200.atari import _Space
from softtreemax_b200.policy import ActorCriticCnnPolicyDepth0
from softtreemax_b200.ppo import PPO


class BanditVecEnv:
    """N envs; observation = a constant image whose brightness names the rewarded action; episodes of 5 steps."""

    def __init__(self, n, n_actions=3):
        self.num_envs, self.A = n, n_actions
        self.observation_space, self.action_space = _Space(shape=(4, 84, 84)), _Space(n=n_actions)
        self.rng = np.random.RandomState(1)

    def _obs(self):
        o = np.zeros((self.num_envs, 4, 84, 84), np.uint8)
        o += (self.target * 100 + 20).astype(np.uint8)[:, None, None, None]
        return o

    def reset(self):
        self.target = self.rng.randint(self.A, size=self.num_envs)
        self.t = np.zeros(self.num_envs, np.int64)
        return self._obs()

    def step(self, a):
        a = np.asarray(a).reshape(-1)
        r = (a == self.target).astype(np.float32)
        self.t += 1
        done = (self.t >= 5).astype(np.float32)
        for i in np.nonzero(done)[0]:
            self.target[i] = self.rng.randint(self.A)
            self.t[i] = 0
        return self._obs(), r, done, [{"orig_reward": float(x)} for x in r]


def test_vectorised_rollout_gae_and_update():
    torch.manual_seed(0)
    env = BanditVecEnv(6)
    ppo = PPO(ActorCriticCnnPolicyDepth0, env, n_steps=8, batch_size=16, n_epochs=2, seed=0, policy_kwargs=dict(device="cpu"))
    obs = env.reset()
    ppo._ep_ret = np.zeros(6)
    buf, obs2, start, ep = ppo._rollout_vec(obs, np.ones(6, np.float32))
    assert buf["obs"].shape == (48, 4, 84, 84) and buf["act"].shape == (48,) and ppo.num_timesteps == 48
    assert len(ep) == 6                                   # every env finished one 5-step episode in 8 steps
    # GAE against a direct per-env recomputation
    n, N = 8, 6
    rew, val, st = (buf[k].reshape(n, N) for k in ("rew", "val", "start"))
    with torch.no_grad():
        last = ppo.policy.forward(torch.from_numpy(obs2))[1].reshape(-1).numpy()
    for e in range(N):
        gae, want = 0.0, np.zeros(n)
        for t in reversed(range(n)):
            nt = 1.0 - (start[e] if t == n - 1 else st[t + 1, e])
            nv = last[e] if t == n - 1 else val[t + 1, e]
            gae = rew[t, e] + 0.99 * nv * nt - val[t, e] + 0.99 * 0.95 * nt * gae
            want[t] = gae
        assert np.allclose(buf["adv"].reshape(n, N)[:, e], want, atol=1e-5)
    before = [p.detach().clone() for p in ppo.policy.parameters()]
    stats = ppo._train(buf, 1.0)
    assert np.all(np.isfinite(stats))
    assert any(not torch.equal(a, b) for a, b in zip(before, ppo.policy.parameters()))
    assert "action_net.weight" in ppo.policy.gradients_history


def test_learn_loop_on_vec_env():
    torch.manual_seed(1)
    ppo = PPO(ActorCriticCnnPolicyDepth0, BanditVecEnv(4), n_steps=4, batch_size=8, n_epochs=1, seed=0,
              policy_kwargs=dict(device="cpu")).learn(32)
    assert ppo.num_timesteps == 32 and len(ppo.log) == 2 and all(np.isfinite(r["loss"]) for r in ppo.log)
    a, _ = ppo.policy.predict(np.zeros((3, 4, 84, 84), np.uint8), deterministic=True)
    assert a.shape == (3,)


def test_training_callback_records_episode_metrics(tmp_path):
    from softtreemax_b200.callbacks import TrainingCallback
    torch.manual_seed(2)
    cb = TrainingCallback(path=str(tmp_path / "train.jsonl"))
    ppo = PPO(ActorCriticCnnPolicyDepth0, BanditVecEnv(2), n_steps=10, batch_size=10, n_epochs=1, seed=0,
              policy_kwargs=dict(device="cpu")).learn(60, callback=cb)
    assert len(cb.records) == 6                                     # env 0 finishes an episode every 5 steps, 30 steps
    assert all(r["train/episodic_length"] == 5 for r in cb.records)
    assert all(0.0 <= r["train/episodic_reward"] <= 5.0 for r in cb.records)
    assert cb.records[-1]["train/policy_weights_grad_var"] > 0.0     # two or more updates have left gradient history
    assert all(len(h) <= 100 for h in ppo.policy.gradients_history.values())
    assert len(open(tmp_path / "train.jsonl").read().splitlines()) == 6
