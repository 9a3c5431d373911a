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


def test_width_limited_tree_matches_oracle_replay():
    """max_width != -1 (cule_bfs.py:81-96,145-150,175-196): the survivor draw is random, so the GPU run's
    survivor indices are replayed through the oracle; leaves, rewards and first_action must then match
    bit for bit, and the policy's logits must match the restatement with first_action."""
    from oracle import orc, policy_ref
    env, pol = make_policy(depth=3, max_width=2)              # 2*A = 8 nodes per level at most
    obs = env.reset()
    for a in [1, 3, 3, 0, 2]:
        obs, *_ = env.step(a)
    bfs = pol.cule_bfs
    A = env.action_space.n
    obs_t = torch.from_numpy(obs).float()[None]
    torch.manual_seed(1)
    with torch.no_grad():
        pol.action_net.weight.mul_(30.0)
    leaves, rewards, first_action = bfs.bfs(obs_t, 3)
    assert leaves.shape == (8, 4, 84, 84) and first_action.shape == (8, 1) and len(bfs.last_survivors) == 2
    # oracle replay
    g = orc.Game("synth_breakout")
    st = env.env.blobs[0].cpu().numpy()[None].copy()
    rw = np.zeros(1, np.float32); dn = np.zeros(1, np.uint8); lv = np.array([g.lives(st[0])], np.int32)
    stacks = obs[None].copy()
    fa = np.arange(A)[:, None]
    pruned = 0
    for depth in range(3):
        n = st.shape[0] * A
        st = np.repeat(st, A, axis=0); rw = np.repeat(rw, A); dn = np.repeat(dn, A); lv = np.repeat(lv, A)
        if depth != 0:
            fa = np.repeat(fa, A, axis=0)
        new = np.zeros((n, 4, 84, 84), np.uint8)
        for i in range(n):
            o, *_ = g.agent_step(st[i], i % A, weight=float(np.float32(0.99) ** depth) if False else 0.99 ** depth,
                                 done=dn[i:i + 1], reward=rw[i:i + 1], lives=lv[i:i + 1])
            new[i, :3] = stacks[i // A, 1:]
            new[i, 3] = o
        stacks = new
        if depth != 2 and n > 2:
            top = np.array(bfs.last_survivors[pruned]); pruned += 1
            fa, stacks, st, rw, dn, lv = fa[top], stacks[top], st[top], rw[top], dn[top], lv[top]
    assert np.array_equal(leaves.cpu().numpy(), stacks)
    assert np.array_equal(rewards.cpu().numpy(), np.sign(rw))
    assert np.array_equal(first_action.cpu().numpy(), fa)
    # policy logits with first_action
    actions, value, log_prob = pol.forward(obs_t, deterministic=True)
    key = int(pol.hash_obs(obs_t.to(pol.device))[0].item())
    leaves2, rewards2, fa2 = pol.obs2leaves_dict[key]
    params = {k: v.cpu() for k, v in pol.kernel_params().items()}
    _, head, _ = policy_ref.nature_cnn_forward(params, leaves2.cpu())
    logits, _, _ = policy_ref.softtreemax_logits(head[None], rewards2.cpu()[None], A, 0.99 ** 3, float(pol.beta), 1.0, 0.001,
                                                 False, first_action=fa2.cpu().reshape(1, -1))
    got = torch.distributions.Categorical(logits=logits).log_prob(actions.cpu())
    assert abs(float(got) - float(log_prob)) <= 5e-2
    # and the update path accepts the grouped leaves
    v, lp, ent = pol.evaluate_actions(obs_t, actions)
    (lp.sum() + v.sum()).backward()
    assert torch.isfinite(pol.beta.grad)


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
    for name in ("synth_breakout", "synth_pong", "synth_pacman", "synth_torture"):      # incl. F8 banking and the undocumented opcodes
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


def test_cule_env_multiple_matches_oracle_and_restarts():
    """Depth-0 arm (cule_env_multiple.py): N root envs stepped together through K2 + K3 == N oracle envs stepped
    one by one, stacks included; a finished env restarts from power-on + one no-op with a zeroed stack."""
    from oracle import orc
    from softtreemax_b200.envs import CuleEnvMultiple
    N = 37
    env = CuleEnvMultiple(device="cuda:0", env_kwargs=ENV_KWARGS, clip_reward=True, n_envs=N)
    env.seed(3)
    noops = np.random.RandomState(3).randint(31)
    stack = env.reset()
    g = orc.Game("synth_breakout")
    s0 = g.reset()
    lives = np.array([g.lives(s0)], np.int32)
    obs = None
    for _ in range(max(1, noops)):
        obs, _, _, _ = g.agent_step(s0, 0, lives=lives)
    want = np.zeros((4, 84, 84), np.uint8)
    want[3] = obs
    assert stack.shape == (N, 4, 84, 84) and all(np.array_equal(stack[i], want) for i in range(N))
    states = [s0.copy() for _ in range(N)]
    stacks = [want.copy() for _ in range(N)]
    lv = [lives.copy() for _ in range(N)]
    fresh = g.reset()
    fl = np.array([g.lives(fresh)], np.int32)
    g.agent_step(fresh, 0, lives=fl)
    ended = np.zeros(N, bool)
    rng = np.random.default_rng(5)
    saw_done = False
    for t in range(260):
        a = rng.integers(0, 4, N)
        a[:5] = 0                                           # a few envs never serve: they lose all lives and restart
        got_stack, got_r, got_d, infos = env.step(a)
        for i in range(N):
            if ended[i]:
                states[i], lv[i], stacks[i] = fresh.copy(), fl.copy(), np.zeros((4, 84, 84), np.uint8)
            o, r, d, l = g.agent_step(states[i], int(a[i]), lives=lv[i])
            stacks[i] = np.concatenate([stacks[i][1:], o[None]])
            ended[i] = bool(d)
            assert np.array_equal(got_stack[i], stacks[i]), (t, i)
            assert got_r[i] == np.sign(r) and infos[i]["orig_reward"] == r and bool(got_d[i]) == bool(d), (t, i)
            assert infos[i]["ale.lives"] == l
        saw_done |= bool(ended.any())
    assert saw_done, "the trace must cover a restart"


def test_depth0_ppo_on_cule_env_multiple(tmp_path):
    from softtreemax_b200.envs import CuleEnvMultiple
    from softtreemax_b200.policy import ActorCriticCnnPolicyDepth0
    from softtreemax_b200.ppo import PPO
    env = CuleEnvMultiple(device="cuda:0", env_kwargs=ENV_KWARGS, clip_reward=True, n_envs=16)
    torch.manual_seed(0)
    model = PPO(ActorCriticCnnPolicyDepth0, env, n_steps=16, batch_size=64, n_epochs=2, seed=0).learn(512)
    assert model.num_timesteps == 512 and len(model.log) == 2 and all(np.isfinite(r["loss"]) for r in model.log)
    # save / load round trip and the evaluation loop of main.py:87-99 (vectorised and single-env)
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.ppo import evaluate_policy
    path = str(tmp_path / "depth0.pt")
    model.policy.save(path)
    pol = ActorCriticCnnPolicyDepth0.load(path)
    for a, b in zip(pol.state_dict().values(), model.policy.state_dict().values()):
        assert torch.equal(a, b)
    scores, lengths = evaluate_policy(pol, env, n_eval_episodes=3, return_episode_rewards=True)
    assert len(scores) == 3 and all(l > 0 for l in lengths)
    single = CuleEnv(device="cuda:0", env_kwargs=ENV_KWARGS, clip_reward=True, fire_reset=True)
    mean, std = evaluate_policy(pol, single, n_eval_episodes=2)
    assert np.isfinite(mean) and np.isfinite(std)


def test_forward_roots_equals_single_root_forwards():
    """forward_roots (R roots through one bfs, one K4 pass, one K5 launch) against R calls of forward(): same logits
    (compared through log_prob of the same action) and values.  K4 tiles images in groups of 128 rows, so a root evaluated
    in a batch of 6 and alone goes through identical arithmetic: the tolerance is exact equality of the logits."""
    import numpy as np
    import torch as th
    from oracle import orc
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    kw = dict(env_name="PongNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    np.random.seed(1)
    th.manual_seed(1)
    env = CuleEnv(device="cuda:0", env_kwargs=kw, clip_reward=True, fire_reset=True)
    pol = ActorCriticCnnTSPolicy(env.observation_space, env.action_space, lambda _: 1e-4, tree_depth=2, gamma=0.99, step_env=env,
                                 buffer_size=128, learn_alpha=True, learn_beta=True, max_width=-1, use_leaves_v=False,
                                 is_cumulative_mode=False, regularization=0.001)
    g = orc.Game("synth_pong")
    roots, stacks = orc.make_roots(g, 6, seed=3, base_steps=50, spread=30)
    blobs, obs = th.from_numpy(roots).cuda(), th.from_numpy(stacks).cuda()
    _, values, _, logits = pol.forward_roots(obs, blobs, deterministic=True)
    for r in range(6):
        env.env.blobs[:1] = blobs[r:r + 1]                      # the root emulator state CuleBFS.bfs injects (cule_bfs.py:110-112)
        a, v, lp = pol.forward(obs[r:r + 1].float(), deterministic=True)
        ref = th.log_softmax(logits[r], 0)
        assert int(a) == int(logits[r].argmax())
        assert abs(float(lp) - float(ref[int(a)])) <= 1e-6
        assert abs(float(v.reshape(-1)[0]) - float(values[r])) <= 1e-6
