"""K4 backward (stm_cnn_forward_train / stm_cnn_backward, tcgen05 data- and weight-gradient GEMMs) against torch autograd of the
same network.  Two references, tolerances stated per parameter tensor (g = kernel gradient):
  (1) autograd through a torch emulation of the kernel's forward roundings (weights and inter-layer activations rounded to
      bf16, conv1 weights to fp16, straight-through) — the same ReLU masks as the kernel:
          ||g - g_emu|| / ||g_emu|| <= 2.5e-2   and   max|g - g_emu| <= 6e-2 * max|g_emu|
      (what is left is the bf16 rounding of the activation gradients between layers);
  (2) plain fp32 autograd (oracle/policy_ref.py's network): ||g - g_fp32|| / ||g_fp32|| <= 0.2 — dominated by ReLU units whose
      pre-activation is within the forward's bf16 error of zero and so take the other branch (each flips its whole gradient).
Image counts are the PPO minibatches of SURVEY Appendix A: Breakout d2 544, Breakout d3 2 080, MsPacman d2 2 624."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
GRAD_ATOL_FRAC, GRAD_REL_L2, GRAD_REL_L2_FP32 = 6e-2, 2.5e-2, 0.2


def torch_reference(params, obs, dheads, emulate):
    """autograd of the network; emulate = with the kernel's forward roundings (straight-through)."""
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    if emulate:
        rb = lambda t: t + (t.to(torch.bfloat16).float() - t).detach()
        rh = lambda t: t + (t.to(torch.float16).float() - t).detach()
    else:
        rb = rh = lambda t: t
    x = obs.float()
    x = rb(F.relu(F.conv2d(x, rh(p["cnn.0.weight"]), None, stride=4) / 255.0 + p["cnn.0.bias"].view(1, -1, 1, 1)))
    x = rb(F.relu(F.conv2d(x, rb(p["cnn.2.weight"]), p["cnn.2.bias"], stride=2)))
    x = rb(F.relu(F.conv2d(x, rb(p["cnn.4.weight"]), p["cnn.4.bias"], stride=1)))
    f = rb(F.relu(F.linear(x.flatten(1), rb(p["linear.0.weight"]), p["linear.0.bias"])))
    heads = torch.cat([F.linear(f, rb(p["action_net.weight"]), p["action_net.bias"]), F.linear(f, rb(p["value_net.weight"]), p["value_net.bias"])], dim=1)
    (heads * dheads).sum().backward()
    return heads.detach(), {k: v.grad for k, v in p.items()}


@pytest.mark.parametrize("n,A", [(37, 4), (544, 4), (2080, 4), (2624, 9)])
def test_cnn_backward_matches_torch_autograd(n, A):
    from oracle import policy_ref
    from softtreemax_b200 import ops
    from softtreemax_b200.engine import Engine
    eng = Engine("synth_breakout")
    try:
        dev = eng.device
        params = {k: v.to(dev) for k, v in policy_ref.make_params(A, seed=4).items()}
        params["action_net.weight"] = params["action_net.weight"] * 30
        g = torch.Generator().manual_seed(n)
        obs = torch.randint(0, 256, (n, 4, 84, 84), generator=g, dtype=torch.uint8)
        obs[:, :, 40:, :] //= 8
        obs = obs.to(dev)
        dheads = torch.randn((n, A + 1), generator=g).to(dev) / n
        eng.cnn_set_weights(params)
        plist = [params[k].clone().requires_grad_(True) for k in eng.PARAM_KEYS]
        heads = ops.cnn_heads(eng, obs, plist)
        (heads * dheads).sum().backward()
        torch.cuda.synchronize()
        emu_heads, emu = torch_reference(params, obs, dheads, emulate=True)
        ref_heads, ref = torch_reference(params, obs, dheads, emulate=False)
        assert (heads.detach() - emu_heads).abs().max() <= 5e-3 * ref_heads.abs().max()
        assert (heads.detach() - ref_heads).abs().max() <= 3e-2 * ref_heads.abs().max()
        bad, report = [], []
        for k, p in zip(eng.PARAM_KEYS, plist):
            gk, ge, gr = p.grad, emu[k], ref[k]
            assert gk is not None and gk.shape == gr.shape and torch.isfinite(gk).all(), k
            scale = ge.abs().max().item()
            err = (gk - ge).abs().max().item()
            rel = ((gk - ge).norm() / ge.norm()).item()
            rel32 = ((gk - gr).norm() / gr.norm()).item()
            report.append(f"{k}: vs emulation max err {err:.3e} (scale {scale:.3e}), rel L2 {rel:.3e}; vs fp32 rel L2 {rel32:.3e}")
            if not (err <= GRAD_ATOL_FRAC * scale and rel <= GRAD_REL_L2 and rel32 <= GRAD_REL_L2_FP32):
                bad.append(k)
        print("\n".join(report))
        assert not bad, "gradient out of tolerance for " + ", ".join(bad) + "\n" + "\n".join(report)
    finally:
        eng.close()


def test_ppo_losses_match_between_kernel_and_torch_update_paths():
    """evaluate_actions with the CNN on K4 (forward_train + tcgen05 backward) against the same policy with the CNN in torch
    fp32 autograd: same rollout, same minibatches, three Adam updates each.  Stated tolerance: every loss term of every update
    |loss_kernel - loss_torch| <= 2e-3 + 2e-2 * |loss_torch| (policy-gradient, value and entropy terms separately), and beta
    (the learnable temperature) after the three updates within 1e-3."""
    import numpy as np
    from softtreemax_b200.envs import CuleEnv
    from softtreemax_b200.policy import ActorCriticCnnTSPolicy
    kw = dict(env_name="BreakoutNoFrameskip-v4", color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True, frameskip=4)
    np.random.seed(7)
    torch.manual_seed(7)
    env = CuleEnv(device="cuda:0", env_kwargs=kw, clip_reward=True, fire_reset=True)
    mk = lambda backend: ActorCriticCnnTSPolicy(env.observation_space, env.action_space, lambda _: 2.5e-4, tree_depth=2, gamma=0.99,
                                                step_env=env, buffer_size=128, learn_alpha=False, learn_beta=True, max_width=-1,
                                                use_leaves_v=False, is_cumulative_mode=False, regularization=0.001,
                                                update_backend=backend)
    pk, pt = mk("kernel"), mk("torch")
    pt.load_state_dict(pk.state_dict())
    B = 24
    obs_l, act_l, lp_l, seen = [], [], [], set()
    obs = env.reset()
    while len(obs_l) < B:
        o = torch.from_numpy(np.asarray(obs)).float()[None]
        with torch.no_grad():
            a, v, lp = pk.forward(o)
            pt.forward(o)                                    # fills pt's leaf cache with the same tree
        key = int(pk.hash_obs(o)[0])
        if key not in seen:
            seen.add(key); obs_l.append(np.asarray(obs).copy()); act_l.append(int(a)); lp_l.append(float(lp))
        obs, _, done, _ = env.step(int(a))
        if done:
            obs = env.reset()
    ob = torch.from_numpy(np.stack(obs_l)).float()
    act = torch.tensor(act_l, device="cuda:0")
    old_lp = torch.tensor(lp_l, device="cuda:0")
    adv = torch.linspace(-1.5, 1.5, B, device="cuda:0")
    ret = torch.linspace(-0.5, 0.5, B, device="cuda:0")

    def update(pol):
        values, logp, ent = pol.evaluate_actions(ob, act)
        ratio = torch.exp(logp - old_lp)
        pg = -torch.min(adv * ratio, adv * torch.clamp(ratio, 0.9, 1.1)).mean()
        vl = F.mse_loss(ret, values.reshape(-1))
        el = -ent.mean()
        pol.optimizer.zero_grad()
        (pg + 0.01 * el + 1.0 * vl).backward()
        torch.nn.utils.clip_grad_norm_(pol.parameters(), 0.5)
        pol.optimizer.step()
        return [float(pg.detach()), float(vl.detach()), float(el.detach())]

    for it in range(3):
        lk, lt = update(pk), update(pt)
        for name, a, b in zip(("policy-gradient", "value", "entropy"), lk, lt):
            assert abs(a - b) <= 2e-3 + 2e-2 * abs(b), f"update {it}, {name} loss: kernel {a:.6f} vs torch {b:.6f}"
    assert abs(float(pk.beta) - float(pt.beta)) <= 1e-3
