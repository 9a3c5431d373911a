"""Autograd-aware Python faces of K4 (NatureCNN + heads, update path) and K5 (logits).  The arithmetic is in libstm_b200.so."""
import ctypes as C

import torch

from ._lib import check, lib


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class _SoftTreeMaxLogits(torch.autograd.Function):
    @staticmethod
    def forward(ctx, eng, head, rewards, depth0, beta, alpha, val_coef, eps, cumulative, group):
        R, L, A = head.shape
        head = head.contiguous().float()
        rewards = rewards.contiguous().float()
        d0 = depth0.contiguous().float() if depth0 is not None else None
        logits = torch.empty((R, A), dtype=torch.float32, device=head.device)
        b = float(beta.detach()) if isinstance(beta, torch.Tensor) else float(beta)
        a = float(alpha.detach()) if isinstance(alpha, torch.Tensor) else float(alpha)
        grp = group.contiguous().to(torch.int32) if group is not None else None
        check(lib().stm_logits_fwd(eng._h, _p(head), _p(rewards), _p(d0), None, _p(grp), R, L, A, float(val_coef), b, a,
                                   float(eps), int(cumulative), _p(logits), None, None, eng._stream()), eng._h)
        ctx.eng, ctx.scal, ctx.grp = eng, (R, L, A, float(val_coef), b, a, float(eps), int(cumulative)), grp
        ctx.save_for_backward(head, rewards, d0 if d0 is not None else torch.empty(0, device=head.device))
        ctx.has_d0 = d0 is not None
        ctx.beta_t = isinstance(beta, torch.Tensor)
        ctx.alpha_t = isinstance(alpha, torch.Tensor)
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        head, rewards, d0 = ctx.saved_tensors
        R, L, A, val_coef, b, a, eps, cum = ctx.scal
        eng = ctx.eng
        d0 = d0 if ctx.has_d0 else None
        dlogits = dlogits.contiguous().float()
        dhead = torch.empty_like(head)
        dd0 = torch.empty((R, A), dtype=torch.float32, device=head.device) if d0 is not None else None
        dbeta = torch.empty((R,), dtype=torch.float32, device=head.device)
        dalpha = torch.empty((R,), dtype=torch.float32, device=head.device)
        check(lib().stm_logits_bwd(eng._h, _p(head), _p(rewards), _p(d0), _p(ctx.grp), R, L, A, val_coef, b, a, eps, cum,
                                   _p(dlogits), _p(dhead), _p(dd0), _p(dbeta), _p(dalpha), eng._stream()), eng._h)
        return (None, dhead, None, dd0, dbeta.sum() if ctx.beta_t else None,
                dalpha.sum() if ctx.alpha_t else None, None, None, None, None)


def softtreemax_logits(eng, head, rewards, val_coef, beta, alpha, eps, cumulative, depth0=None, group=None):
    """head (R, L, A), rewards (R, L) -> regularised logits (R, A); differentiable in head, depth0,
    beta and alpha (tensors).  group (R, L) int: root action of every leaf of a width-limited tree.
    policies/actor_critic_ts.py:60-95 / :147-188."""
    return _SoftTreeMaxLogits.apply(eng, head, rewards, depth0, beta, alpha, val_coef, eps, cumulative, group)


def softtreemax_logits_full(eng, head, rewards, val_coef, beta, alpha, eps, cumulative, depth0=None, values=None,
                            group=None):
    """No-grad variant that also returns the un-mixed logits and the use_leaves_v root value."""
    R, L, A = head.shape
    head, rewards = head.contiguous().float(), rewards.contiguous().float()
    d0 = depth0.contiguous().float() if depth0 is not None else None
    v = values.contiguous().float() if values is not None else None
    logits = torch.empty((R, A), dtype=torch.float32, device=head.device)
    raw = torch.empty_like(logits)
    vmax = torch.empty((R,), dtype=torch.float32, device=head.device) if v is not None else None
    grp = group.contiguous().to(torch.int32) if group is not None else None
    check(lib().stm_logits_fwd(eng._h, _p(head), _p(rewards), _p(d0), _p(v), _p(grp), R, L, A, float(val_coef), float(beta),
                               float(alpha), float(eps), int(cumulative), _p(logits), _p(raw), _p(vmax),
                               eng._stream()), eng._h)
    return logits, raw, vmax


TRAIN_CHUNK = 32768          # images per stm_cnn_forward_train / stm_cnn_backward pair


class _CnnHeads(torch.autograd.Function):
    """heads (n, A+1) = NatureCNN + action_net / value_net on u8 observation stacks, differentiable in the twelve
    parameters: forward = stm_cnn_forward_train (activations kept in the ctx), backward = stm_cnn_backward.  A batch larger
    than TRAIN_CHUNK images is evaluated in chunks and its backward re-runs the forward of each chunk (the ctx keeps the
    activations of one chunk); so does a backward that finds the ctx's activations replaced by a later forward_train."""

    @staticmethod
    def forward(ctx, eng, obs_u8, *params):
        ctx.eng, ctx.obs, ctx.params = eng, obs_u8, params
        n = obs_u8.shape[0]
        if n <= TRAIN_CHUNK:
            eng._train_token = ctx.token = object()
            return eng.cnn_forward_train(obs_u8)
        ctx.token = None
        head, val = eng.cnn_forward(obs_u8)
        return torch.cat([head, val[:, None]], dim=1)

    @staticmethod
    def backward(ctx, dheads):
        eng, obs = ctx.eng, ctx.obs
        dheads = dheads.contiguous().float()
        total = None
        for i0 in range(0, obs.shape[0], TRAIN_CHUNK):
            o, d = obs[i0:i0 + TRAIN_CHUNK], dheads[i0:i0 + TRAIN_CHUNK]
            if ctx.token is None or getattr(eng, "_train_token", None) is not ctx.token:
                eng.cnn_forward_train(o)                                  # recompute this chunk's activations
                eng._train_token = None
            g = eng.cnn_backward(o, d, ctx.params)
            total = g if total is None else [a + b for a, b in zip(total, g)]
        return (None, None) + tuple(g.to(p.dtype) for g, p in zip(total, ctx.params))


def cnn_heads(eng, obs_u8, params):
    """obs (n, 4, 84, 84) uint8 on the device, params: the twelve tensors of Engine.PARAM_KEYS (in that order; the Engine must
    hold their current values: Engine.cnn_set_weights) -> heads (n, A+1): columns 0..A-1 action_net, column A value_net.
    Replaces extract_features -> action_net / value_net under autograd (actor_critic_ts.py:139-149)."""
    return _CnnHeads.apply(eng, obs_u8.contiguous(), *params)
