"""K4 (NatureCNN + heads on tcgen05).  The oracle is a torch fp32 restatement of SB3's NatureCNN
(oracle/policy_ref.py).  Tolerances, stated:
  - against the SAME network evaluated in torch with bf16-rounded weights (fp16-rounded for conv1) and bf16-rounded
    inter-layer activations (what the kernel computes, fp32 accumulation): |err| <= 5e-3 * scale (a bf16 rounding tie in one activation shifts its successors);
  - against the plain fp32 network: |err| <= 3e-2 * scale  (bf16 has 8 mantissa bits),
where scale = max |reference output| over the batch."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def bf16_emulation(params, obs_u8):
    r = lambda t: t.to(torch.bfloat16).float()
    h = lambda t: t.to(torch.float16).float()          # conv1 runs on fp16 operands (u8 pixels are exact in fp16)
    x = obs_u8.float()
    x = r(F.relu(F.conv2d(x, h(params["cnn.0.weight"]), None, stride=4) / 255.0 + params["cnn.0.bias"].view(1, -1, 1, 1)))
    x = r(F.relu(F.conv2d(x, r(params["cnn.2.weight"]), params["cnn.2.bias"], stride=2)))
    x = r(F.relu(F.conv2d(x, r(params["cnn.4.weight"]), params["cnn.4.bias"], stride=1)))
    f = r(F.relu(F.linear(x.flatten(1), r(params["linear.0.weight"]), params["linear.0.bias"])))
    return F.linear(f, r(params["action_net.weight"]), params["action_net.bias"]), \
        F.linear(f, r(params["value_net.weight"]), params["value_net.bias"])


@pytest.mark.parametrize("n,A", [(1, 4), (5, 4), (130, 6), (700, 9), (33, 18)])
def test_cnn_forward_matches_torch(n, A):
    from oracle import policy_ref
    from softtreemax_b200.engine import Engine
    eng = Engine("synth_breakout")
    try:
        params = policy_ref.make_params(A, seed=4)
        # make the heads non-trivial (SB3's 0.01-gain policy head would hide errors)
        params["action_net.weight"] = params["action_net.weight"] * 50
        g = torch.Generator().manual_seed(n)
        obs = torch.randint(0, 256, (n, 4, 84, 84), generator=g, dtype=torch.uint8)
        obs[:, :, 40:, :] //= 8                                     # some structure: dim lower half
        eng.cnn_set_weights(params)
        act, val = eng.cnn_forward(obs.to(eng.device))
        torch.cuda.synchronize()
        _, ref_a, ref_v = policy_ref.nature_cnn_forward(params, obs)
        emu_a, emu_v = bf16_emulation(params, obs)
        sa, sv = ref_a.abs().max().item(), ref_v.abs().max().item()
        assert torch.isfinite(act).all() and torch.isfinite(val).all()
        assert (act.cpu() - emu_a).abs().max().item() <= 5e-3 * sa, "vs bf16 emulation (action head)"
        assert (val.cpu() - emu_v[:, 0]).abs().max().item() <= 5e-3 * sv, "vs bf16 emulation (value)"
        assert (act.cpu() - ref_a).abs().max().item() <= 3e-2 * sa, "vs fp32 reference (action head)"
        assert (val.cpu() - ref_v[:, 0]).abs().max().item() <= 3e-2 * sv, "vs fp32 reference (value)"
    finally:
        eng.close()


def test_cnn_on_real_leaf_stacks(orc=None):
    """Observation stacks produced by stm_bfs go straight into stm_cnn_forward (u8, no conversion)."""
    from oracle import orc as _orc, policy_ref
    from softtreemax_b200.engine import Engine
    g = _orc.Game("synth_breakout")
    eng = Engine("synth_breakout")
    try:
        roots, stacks = _orc.make_roots(g, 6, seed=8, base_steps=30, spread=9)
        out = eng.bfs(torch.from_numpy(roots).to(eng.device), torch.from_numpy(stacks).to(eng.device), 2)
        params = policy_ref.make_params(4, seed=1)
        eng.cnn_set_weights(params)
        act, val = eng.cnn_forward(out["stacks"])
        torch.cuda.synchronize()
        _, ref_a, ref_v = policy_ref.nature_cnn_forward(params, out["stacks"].cpu())
        assert (act.cpu() - ref_a).abs().max().item() <= 3e-2 * ref_a.abs().max().item() + 1e-6
        assert (val.cpu() - ref_v[:, 0]).abs().max().item() <= 3e-2 * ref_v.abs().max().item()
    finally:
        eng.close()


def test_cnn_crosses_the_chunk_boundary():
    """stm_cnn_forward works in chunks of 32 768 images (csrc/stm_cnn.cu); n = 32 768 + 33 crosses that boundary.
    A random sample of 256 rows, always including the rows either side of the boundary and the very last one, is
    compared against the fp32 restatement and the bf16 emulation with the tolerances stated at the top."""
    from oracle import policy_ref
    from softtreemax_b200.engine import Engine
    n, A = 32768 + 33, 4
    eng = Engine("synth_breakout")
    try:
        params = policy_ref.make_params(A, seed=9)
        params["action_net.weight"] = params["action_net.weight"] * 50
        g = torch.Generator().manual_seed(77)
        base = torch.randint(0, 256, (512, 4, 84, 84), generator=g, dtype=torch.uint8)
        idx = torch.randint(0, 512, (n,), generator=g)
        obs = base[idx].contiguous()
        obs[:, 0, 0, 0] = (torch.arange(n) % 251).to(torch.uint8)                     # every image differs
        obs[:, 3, 83, 83] = (torch.arange(n) // 251 % 251).to(torch.uint8)
        eng.cnn_set_weights(params)
        act, val = eng.cnn_forward(obs.to(eng.device))
        torch.cuda.synchronize()
        pick = torch.cat([torch.tensor([0, 32766, 32767, 32768, 32769, n - 2, n - 1]),
                          torch.randint(0, n, (249,), generator=g)])
        _, ref_a, ref_v = policy_ref.nature_cnn_forward(params, obs[pick])
        emu_a, emu_v = bf16_emulation(params, obs[pick])
        sa, sv = ref_a.abs().max().item(), ref_v.abs().max().item()
        assert torch.isfinite(act).all() and torch.isfinite(val).all()
        assert (act[pick].cpu() - emu_a).abs().max().item() <= 5e-3 * sa
        assert (val[pick].cpu() - emu_v[:, 0]).abs().max().item() <= 5e-3 * sv
        assert (act[pick].cpu() - ref_a).abs().max().item() <= 3e-2 * sa
        assert (val[pick].cpu() - ref_v[:, 0]).abs().max().item() <= 3e-2 * sv
    finally:
        eng.close()


# Tolerance of the SoftTreeMax logits (K4 -> K5) against the fp32 restatement of the whole policy
# (oracle/policy_ref.py: NatureCNN in fp32 -> actor_critic_ts.py:60-95 -> utils.py:52-57), per element:
#     |logit_kernel - logit_fp32| <= LOGIT_ATOL + LOGIT_RTOL * |logit_fp32|
# on SB3-initialised weights (policy-head gain 0.01) and with the head scaled x50 (a trained, peaked policy).
LOGIT_ATOL, LOGIT_RTOL = 4e-3, 1e-2


@pytest.mark.parametrize("name,A,depth", [("Breakout d2", 4, 2), ("Breakout d3", 4, 3), ("Pong d3", 6, 3), ("MsPacman d2", 9, 2)])
@pytest.mark.parametrize("head_gain", [1.0, 50.0])
def test_softtreemax_logits_tolerance(name, A, depth, head_gain):
    from oracle import policy_ref
    from softtreemax_b200 import ops
    from softtreemax_b200.engine import Engine
    R, L = 3, A ** depth
    eng = Engine("synth_breakout")
    try:
        params = policy_ref.make_params(A, seed=4)
        params["action_net.weight"] = params["action_net.weight"] * head_gain
        g = torch.Generator().manual_seed(A * 10 + depth)
        obs = torch.randint(0, 256, (R * L, 4, 84, 84), generator=g, dtype=torch.uint8)
        obs[:, :, 30:, :] //= 6
        rewards = torch.randint(-1, 2, (R, L), generator=g).float()
        eng.cnn_set_weights(params)
        head, val = eng.cnn_forward(obs.to(eng.device))
        for cumulative in (False, True):
            logits, _, vmax = ops.softtreemax_logits_full(eng, head.reshape(R, L, A), rewards.to(eng.device), 0.99 ** depth, 1.0, 1.0,
                                                          0.001, cumulative, values=val.reshape(R, L))
            torch.cuda.synchronize()
            _, ref_head, ref_val = policy_ref.nature_cnn_forward(params, obs)
            ref_logits, _, ref_vmax = policy_ref.softtreemax_logits(ref_head.reshape(R, L, A), rewards, A, 0.99 ** depth, 1.0, 1.0, 0.001,
                                                                    cumulative, values=ref_val.reshape(R, L))
            err = (logits.cpu() - ref_logits).abs()
            bound = LOGIT_ATOL + LOGIT_RTOL * ref_logits.abs()
            assert (err <= bound).all(), f"{name} cumulative={cumulative}: max err {err.max().item():.3e} (bound {bound.min().item():.3e})"
            verr = (vmax.cpu() - ref_vmax).abs()
            assert (verr <= 2e-2 + 2e-2 * ref_vmax.abs()).all(), f"{name}: leaf-value max err {verr.max().item():.3e}"
    finally:
        eng.close()
