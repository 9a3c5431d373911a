"""Generates tests/golden/logits_kat.json — known-answer vectors for K5 (SoftTreeMax logits).

Run in the build container only: it imports the REFERENCE's own `add_regularization_logits`
(/root/reference/utils.py:52-57, the one hot-path function of the reference that imports here) and
restates policies/actor_critic_ts.py:60,66-69,80-86,94-95 line by line in torch around it.  The
vectors travel to the GPU box as a fixture; /root/reference does not.

    python tools/gen_logits_golden.py
"""
import json
import math
import os
import sys

import numpy as np
import torch as th

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_logits(head, rewards, A, depth, gamma, beta, alpha, eps, cumulative, depth0=None, first_action=None):
    """head (L, A), rewards (L,) for ONE root — actor_critic_ts.py forward(); first_action (L, 1) selects
    the max_width != -1 branch (:70-79)."""
    sys.path.insert(0, "/root/reference")
    from utils import add_regularization_logits          # the reference's own function
    val_coef = gamma ** depth                                                      # :60
    mean_actions = val_coef * head + rewards.reshape([-1, 1])                      # :66
    if first_action is None:
        mean_actions_per_subtree = beta * mean_actions.reshape([A, -1])            # :68
        counts = th.ones([1, A]) * mean_actions_per_subtree.shape[1]               # :69
    else:
        mean_actions_per_subtree = th.zeros(A, mean_actions.shape[0], mean_actions.shape[1])     # :71
        idxes = th.arange(mean_actions.shape[0])                                   # :73
        counts = th.zeros(A)                                                       # :74
        v, c = th.unique(first_action, return_counts=True)                         # :75
        counts[v] = c.type(th.float32) * A                                         # :76
        mean_actions_per_subtree[first_action.flatten(), idxes, :] = mean_actions  # :77
        mean_actions_per_subtree = beta * mean_actions_per_subtree.reshape([A, -1])  # :78
        counts = counts.reshape([1, -1])                                           # :79
    if cumulative:
        logits = th.sum(mean_actions_per_subtree, dim=1, keepdim=True).transpose(1, 0) / counts          # :81
    else:
        logits = th.logsumexp(mean_actions_per_subtree, dim=1, keepdim=True).transpose(1, 0) - th.log(counts)  # :84-85
    logits[counts == 0] = -math.inf                                                # :86
    raw = logits.clone()
    d0 = depth0 if depth0 is not None else th.tensor(0)                            # :87
    logits = alpha * logits + (1 - alpha) * d0                                     # :94
    return raw, add_regularization_logits(logits, eps)                             # :95


def survey_inputs(A, d):
    L = A ** d
    l = np.arange(L)[:, None]
    a = np.arange(A)[None, :]
    head = (2.0 * np.sin(0.37 * (l * A + a))).astype(np.float32)
    rew = np.sign(np.cos(1.3 * np.arange(L))).astype(np.float32)
    return th.from_numpy(head), th.from_numpy(rew)


def main():
    cases = []
    # the SURVEY.md §8(c) known-answer configurations
    for A, d, cum in [(4, 2, False), (4, 2, True), (4, 3, False), (6, 3, False), (9, 2, False)]:
        head, rew = survey_inputs(A, d)
        raw, reg = reference_logits(head, rew, A, d, 0.99, 1.0, 1.0, 0.001, cum)
        cases.append(dict(kind="survey", A=A, d=d, cumulative=cum, gamma=0.99, beta=1.0, alpha=1.0, eps=0.001,
                          raw=raw.flatten().tolist(), reg=reg.flatten().tolist()))
    # seeded random cases with gradients (torch autograd through the restatement)
    g = th.Generator().manual_seed(0)
    for A, d, cum, beta, alpha in [(4, 2, False, 1.7, 0.6), (4, 2, True, 0.8, 0.5), (3, 3, False, 2.5, 1.0),
                                   (18, 1, False, 1.0, 0.3), (6, 2, True, 1.3, 1.0)]:
        L = A ** d
        head = (3 * th.randn(L, A, generator=g)).requires_grad_()
        rew = th.randint(-2, 3, (L,), generator=g).float()
        d0 = th.randn(1, A, generator=g).requires_grad_()
        bt = th.tensor(beta, requires_grad=True)
        al = th.tensor(alpha, requires_grad=True)
        raw, reg = reference_logits(head, rew, A, d, 0.97, bt, al, 0.01, cum, d0)
        dy = th.randn(1, A, generator=g)
        (reg * dy).sum().backward()
        cases.append(dict(kind="random", A=A, d=d, cumulative=cum, gamma=0.97, beta=beta, alpha=alpha, eps=0.01,
                          head=head.detach().flatten().tolist(), rewards=rew.tolist(), depth0=d0.detach().flatten().tolist(),
                          raw=raw.detach().flatten().tolist(), reg=reg.detach().flatten().tolist(), dy=dy.flatten().tolist(),
                          dhead=head.grad.flatten().tolist(), ddepth0=d0.grad.flatten().tolist(),
                          dbeta=float(bt.grad), dalpha=float(al.grad)))
    # width-limited trees: explicit first_action per leaf, incl. an action with no surviving leaf
    for A, n, cum, beta in [(4, 12, False, 1.4), (4, 12, True, 0.9), (6, 10, False, 2.0), (9, 27, False, 1.0)]:
        head = (2 * th.randn(n, A, generator=g)).requires_grad_()
        rew = th.randint(-2, 3, (n,), generator=g).float()
        fa = th.randint(0, A - 1, (n, 1), generator=g)             # action A-1 never survives -> -inf
        bt = th.tensor(beta, requires_grad=True)
        raw, reg = reference_logits(head, rew, A, 2, 0.97, bt, 1.0, 0.01, cum, None, fa)
        dy = th.randn(1, A, generator=g)
        (reg * dy).sum().backward()
        cases.append(dict(kind="width", A=A, d=2, n=n, cumulative=cum, gamma=0.97, beta=beta, alpha=1.0, eps=0.01,
                          head=head.detach().flatten().tolist(), rewards=rew.tolist(), first_action=fa.flatten().tolist(),
                          raw=[(v if math.isfinite(v) else None) for v in raw.detach().flatten().tolist()],
                          reg=reg.detach().flatten().tolist(), dy=dy.flatten().tolist(),
                          dhead=head.grad.flatten().tolist(), dbeta=float(bt.grad)))
    out = os.path.join(ROOT, "tests", "golden", "logits_kat.json")
    json.dump(cases, open(out, "w"))
    for c in cases[:5]:
        print(c["A"], c["d"], c["cumulative"], [round(v, 7) for v in c["raw"]], [round(v, 7) for v in c["reg"]])


if __name__ == "__main__":
    main()
