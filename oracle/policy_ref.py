"""CPU restatement (torch fp32) of the policy-side math of the hot path.  TEST INFRASTRUCTURE.

  softtreemax_logits  — policies/actor_critic_ts.py:60,66-69,80-95 batched over roots, + utils.py:52-57
  nature_cnn_forward  — SB3 1.4.0 NatureCNN + heads as reached through extract_features
                        (actor_critic_ts.py:199-218); SB3 is a third-party dependency that is not in
                        /root/reference (pinned `stable-baselines3==1.4.0`, reference Dockerfile:10), so
                        its published architecture is restated: x/255 -> Conv(4,32,8,s4) ReLU ->
                        Conv(32,64,4,s2) ReLU -> Conv(64,64,3,s1) ReLU -> Flatten -> Linear(3136,512) ReLU
                        -> action_net Linear(512,A), value_net Linear(512,1).

Pinned by tests/golden/logits_kat.json (generated with the reference's own add_regularization_logits).
"""
import math

import torch as th
import torch.nn.functional as F


def softtreemax_logits(head, rewards, A, val_coef, beta, alpha, eps, cumulative, depth0=None, values=None,
                       first_action=None):
    """head (R, L, A), rewards (R, L) -> (logits (R, A), raw (R, A), vmax (R,) or None).
    first_action (R, L) int: width-limited tree (actor_critic_ts.py:70-79), else full width (:68-69)."""
    R, L, _ = head.shape
    m = val_coef * head + rewards.unsqueeze(-1)                       # actor_critic_ts.py:66
    if first_action is None:
        per = beta * m.reshape(R, A, -1)                              # :68  row a = leaves under root action a
        cnt = th.full((R, A), float(per.shape[2]))
    else:
        # :71-77 — scatter into a ZERO tensor (A, L, A); non-member slots stay 0 and take part in logsumexp
        sub = th.zeros(R, A, L, A)
        cnt = th.zeros(R, A)
        for r in range(R):
            fa = first_action[r].long()
            sub[r, fa, th.arange(L), :] = m[r]
            v, c = th.unique(fa, return_counts=True)
            cnt[r, v] = c.float() * A
        per = beta * sub.reshape(R, A, -1)
    if cumulative:
        raw = per.sum(dim=2) / cnt                                    # :81
    else:
        raw = th.logsumexp(per, dim=2) - th.log(cnt)                  # :84-85
    raw = th.where(cnt == 0, th.full_like(raw, -math.inf), raw)       # :86
    z = th.where(th.isnan(raw), th.zeros_like(raw), raw)              # :88-90
    d0 = depth0 if depth0 is not None else 0.0
    x = alpha * z + (1 - alpha) * d0                                  # :94
    probs = th.exp(x - th.logsumexp(x, dim=1, keepdim=True))          # utils.py:54
    logits = th.log((1 - eps) * probs + eps / A)                      # utils.py:55-56
    vmax = None
    if values is not None:
        vmax = (val_coef * values + rewards).max(dim=1).values        # :62-63
    return logits, raw, vmax


def nature_cnn_forward(params, obs_u8):
    """params: dict of fp32 tensors (SB3 state_dict naming); obs (N, 4, 84, 84) uint8.
    -> (features (N, 512), action head (N, A), value (N, 1))"""
    x = obs_u8.float() / 255.0
    x = F.relu(F.conv2d(x, params["cnn.0.weight"], params["cnn.0.bias"], stride=4))
    x = F.relu(F.conv2d(x, params["cnn.2.weight"], params["cnn.2.bias"], stride=2))
    x = F.relu(F.conv2d(x, params["cnn.4.weight"], params["cnn.4.bias"], stride=1))
    x = x.flatten(1)
    feat = F.relu(F.linear(x, params["linear.0.weight"], params["linear.0.bias"]))
    return feat, F.linear(feat, params["action_net.weight"], params["action_net.bias"]), \
        F.linear(feat, params["value_net.weight"], params["value_net.bias"])


def make_params(A, seed=4):
    """SB3-style orthogonal init (gain sqrt(2) for the extractor, 0.01 policy head, 1 value head)."""
    g = th.Generator().manual_seed(seed)
    shapes = {"cnn.0": (32, 4, 8, 8), "cnn.2": (64, 32, 4, 4), "cnn.4": (64, 64, 3, 3), "linear.0": (512, 3136),
              "action_net": (A, 512), "value_net": (1, 512)}
    gains = {"action_net": 0.01, "value_net": 1.0}
    p = {}
    for k, shp in shapes.items():
        w = th.empty(shp)
        flat = th.randn(shp[0], int(th.tensor(shp[1:]).prod()), generator=g)
        q, r = th.linalg.qr(flat.t() if flat.shape[0] < flat.shape[1] else flat)
        q = q * th.sign(th.diagonal(r)).unsqueeze(0)
        q = q.t() if flat.shape[0] < flat.shape[1] else q
        w.copy_(q.reshape(shp) * gains.get(k, math.sqrt(2)))
        p[k + ".weight"] = w
        p[k + ".bias"] = th.zeros(shp[0]) if k != "cnn.0" else 0.01 * th.randn(shp[0], generator=g)
    # non-zero biases everywhere so that a dropped bias shows up in parity tests
    for k in list(p):
        if k.endswith(".bias"):
            p[k] = 0.05 * th.randn(p[k].shape, generator=g)
    return p
