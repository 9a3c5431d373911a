"""The drop-in layer: CuleBFS / ActorCriticCnnTSPolicy mirrors over the CUDA path, checked against
the oracle pipeline (oracle bfs -> torch fp32 NatureCNN -> torch logits restatement)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ENV_KWARGS = dict(env_name="BreakoutNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True,
                  episodic_life=True, frameskip=4)                      # main.py:37-38


def make_policy(depth=2, **kw):
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    env = CuleEnv(device="cuda:0", env_kwargs=ENV_KWARGS, clip_reward=True, fire_reset=True)
    args = dict(tree_depth=depth, gamma=0.99, step_env=env, buffer_size=128, learn_alpha=False, learn_beta=True,
                max_width=-1, use_leaves_v=False, is_cumulative_mode=False, regularization=0.001)
    args.update(kw)
    torch.manual_seed(4)
    pol = ActorCriticCnnTSPolicy(env.observation_space, env.action_space, lambda _: 2.5e-4, **args)
    return env, pol


def test_cule_bfs_interface_and_oracle_parity():
    from oracle import orc
    from softtreemax_b200.cule_bfs import CuleBFS
    from softtreemax_b200.envs import CuleEnv
    env = CuleEnv(device="cuda:0", env_kwargs=ENV_KWARGS)
    obs = env.reset()
    for a in [1, 3, 3, 2, 0, 1, 3]:
        obs, r, done, info = env.step(a)
    bfs = CuleBFS(env, tree_depth=2, gamma=0.99)
    assert (bfs.gamma, bfs.max_depth, bfs.max_width) == (0.99, 2, -1)
    leaves, rewards, first_action = bfs.bfs(torch.from_numpy(obs).float()[None], 2)
    assert leaves.shape == (16, 4, 84, 84) and rewards.shape == (16,) and first_action is None
    g = orc.Game("synth_breakout")
    blob = env.env.blobs[0].cpu().numpy()
    want = g.bfs(blob[None], obs[None], 2)
    assert np.array_equal(leaves.cpu().numpy(), want[0][0])
    assert np.array_equal(rewards.cpu().numpy(), want[1][0])
    with pytest.raises(NotImplementedError):
        CuleBFS(env, 2, max_width=4)


@pytest.mark.parametrize("kw", [dict(), dict(is_cumulative_mode=True), dict(learn_alpha=True), dict(use_leaves_v=True),
                                dict(depth=1), dict(depth=3)])
def test_forward_matches_oracle_pipeline(kw):
    from oracle import orc, policy_ref
    env, pol = make_policy(**kw)
    obs = env.reset()
    for a in [1, 3, 0, 2, 2]:
        obs, *_ = env.step(a)
    with torch.no_grad():
        pol.action_net.weight.mul_(40.0)                                 # make the logits spread out
    obs_t = torch.from_numpy(obs).float()[None]
    actions, value, log_prob = pol.forward(obs_t, deterministic=True)
    d = pol.tree_depth
    A = env.action_space.n
    g = orc.Game("synth_breakout")
    stack, rew, *_ = g.bfs(env.env.blobs[0].cpu().numpy()[None], obs[None], d)
    params = {k: v.cpu() for k, v in pol.kernel_params().items()}
    _, head, val = policy_ref.nature_cnn_forward(params, torch.from_numpy(stack[0]))
    _, head0, val0 = policy_ref.nature_cnn_forward(params, torch.from_numpy(obs)[None])
    logits, _, vmax = policy_ref.softtreemax_logits(
        head[None], torch.from_numpy(rew), A, 0.99 ** d, float(pol.beta), float(pol.alpha), 0.001,
        pol.is_cumulative_mode, depth0=head0 if pol.learn_alpha else None, values=val.reshape(1, -1))
    # logits differ only by the bf16 CNN: tolerance 5e-2 absolute on log-probabilities of O(1)
    leaves, rewards, _ = pol.obs2leaves_dict[int(pol.hash_obs(obs_t.to(pol.device))[0].item())]
    assert np.array_equal(leaves.cpu().numpy(), stack[0]) and np.array_equal(rewards.cpu().numpy(), rew[0])
    got = torch.distributions.Categorical(logits=logits).log_prob(actions.cpu())
    assert abs(float(got) - float(log_prob)) <= 5e-2
    want_v = float(vmax) if pol.use_leaves_v else float(val0)
    assert abs(float(value.reshape(-1)[0]) - want_v) <= 3e-2 * max(1.0, abs(want_v))
    # the cache is hit on a repeated observation (actor_critic_ts.py:51-53)
    n = pol.eng.launch_count
    pol.forward(obs_t)
    assert pol.eng.launch_count - n < 20                                # CNN + logits only, no bfs kernels


def test_rollout_then_evaluate_actions_and_checkpoint(tmp_path):
    env, pol = make_policy(depth=2, learn_alpha=True)
    obs = env.reset()
    obs_l, act_l = [], []
    for t in range(6):
        a, v, lp = pol.forward(torch.from_numpy(obs).float()[None])
        obs_l.append(obs)
        act_l.append(int(a))
        obs, r, done, info = env.step(int(a))
        if done:
            obs = env.reset()
    batch = torch.from_numpy(np.stack(obs_l)).float()
    values, log_prob, entropy = pol.evaluate_actions(batch, torch.tensor(act_l))
    assert values.shape == (6, 1) and log_prob.shape == (6,) and entropy.shape == (6,)
    loss = -(log_prob.mean()) + values.pow(2).mean() - 0.01 * entropy.mean()
    pol.optimizer.zero_grad()
    loss.backward()
    assert pol.beta.grad is not None and pol.alpha.grad is not None and torch.isfinite(pol.beta.grad)
    assert pol.features_extractor.cnn[0].weight.grad.abs().sum() > 0
    pol.optimizer.step()
    pol.evaluate_actions(batch, torch.tensor(act_l))
    assert "action_net.weight" in pol.gradients_history              # actor_critic_depth0.py:27-35
    with pytest.raises(KeyError):
        pol.evaluate_actions(batch + 1.0, torch.tensor(act_l))
    path = str(tmp_path / "agent")
    pol.save(path)
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    again = ActorCriticCnnTSPolicy.load(path, env=env, lr_schedule=lambda _: 1e-4)
    for k, v in pol.state_dict().items():
        assert torch.equal(v, again.state_dict()[k])
    a, _ = again.predict(obs[None])
    assert a.shape == (1,)


def test_device_roots_match_oracle_roots():
    """softtreemax_b200.roots.make_roots (CUDA path, long single-env traces with resets) reproduces
    oracle.orc.make_roots bit for bit — 250+ consecutive agent steps of parity."""
    from oracle import orc
    from softtreemax_b200.engine import Engine
    from softtreemax_b200.roots import make_roots
    for name in ("synth_breakout", "synth_pong"):
        eng = Engine(name)
        try:
            st, sk = make_roots(eng, 200, seed=1234)
            want_s, want_k = orc.make_roots(orc.Game(name), 200, seed=1234)
            assert np.array_equal(st.cpu().numpy(), want_s), name
            assert np.array_equal(sk.cpu().numpy(), want_k), name
        finally:
            eng.close()


def test_ppo_loop_runs_and_updates():
    """Two PPO iterations (2 x 128 env steps, 3 epochs x 4 minibatches each) through the mirrored
    policy: rollouts use K1-K5, the update uses K5's backward + torch autograd for the CNN."""
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    from softtreemax_b200.ppo import PPO, get_linear_fn
    env = CuleEnv(device="cuda:0", env_kwargs=ENV_KWARGS, clip_reward=True, fire_reset=True)
    kw = dict(step_env=env, gamma=0.99, tree_depth=1, buffer_size=128, learn_alpha=False, learn_beta=True, max_width=-1,
              use_leaves_v=False, is_cumulative_mode=False, regularization=0.001)
    torch.manual_seed(0)
    model = PPO(ActorCriticCnnTSPolicy, env, learning_rate=get_linear_fn(2.5e-4, 0, 1), clip_range=get_linear_fn(0.1, 0, 1),
                policy_kwargs=kw, seed=0)
    w0 = model.policy.action_net.weight.detach().clone()
    b0 = float(model.policy.beta)
    model.learn(256)
    assert len(model.log) == 2 and all(np.isfinite(r["loss"]) for r in model.log)
    assert not torch.equal(w0, model.policy.action_net.weight.detach())
    assert float(model.policy.beta) != b0
    assert len(model.policy.obs2leaves_dict) <= 128 + 1            # leaf cache is evicted (actor_critic_ts.py:99-103)
