"""Autograd-aware Python faces of K5 (logits).  The arithmetic is in libstm_b200.so."""
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
