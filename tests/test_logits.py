"""K5 (SoftTreeMax logits).  CPU: the torch restatement (oracle/policy_ref.py) against the golden
vectors made with the reference's own add_regularization_logits (tools/gen_logits_golden.py).
GPU: the CUDA kernels (through the C ABI) against the golden vectors and the restatement.
Tolerance: fp32, |err| <= 2e-5 absolute on logits (values are O(1)), 1e-4 relative on gradients."""
import json
import os

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = json.load(open(os.path.join(ROOT, "tests", "golden", "logits_kat.json")))
ATOL = 2e-5


def case_inputs(c):
    A, d = c["A"], c["d"]
    L = c.get("n", A ** d)
    if c["kind"] == "survey":
        l = np.arange(L)[:, None]
        a = np.arange(A)[None, :]
        head = torch.from_numpy((2.0 * np.sin(0.37 * (l * A + a))).astype(np.float32))
        rew = torch.from_numpy(np.sign(np.cos(1.3 * np.arange(L))).astype(np.float32))
        d0 = None
    else:
        head = torch.tensor(c["head"]).reshape(L, A)
        rew = torch.tensor(c["rewards"])
        d0 = torch.tensor(c["depth0"]).reshape(1, A) if "depth0" in c else None
    return head, rew, d0


def case_group(c):
    return torch.tensor(c["first_action"], dtype=torch.int32)[None] if c["kind"] == "width" else None


def want_raw(c):
    return np.array([v if v is not None else -np.inf for v in c["raw"]], np.float32)


@pytest.mark.parametrize("i", range(len(CASES)))
def test_restatement_matches_reference_vectors(i):
    from oracle import policy_ref
    c = CASES[i]
    head, rew, d0 = case_inputs(c)
    if c["kind"] == "width":
        d0 = None
    logits, raw, _ = policy_ref.softtreemax_logits(head[None], rew[None], c["A"], c["gamma"] ** c["d"], c["beta"],
                                                   c["alpha"], c["eps"], c["cumulative"], d0, first_action=case_group(c))
    assert np.allclose(raw[0].numpy(), want_raw(c), atol=ATOL)
    assert np.allclose(logits[0].numpy(), c["reg"], atol=ATOL)


def test_survey_vectors_are_the_published_ones():
    """SURVEY.md §8(c) quotes these numbers; the fixture must reproduce them."""
    c = CASES[0]
    assert np.allclose(c["raw"], [1.7601037, 1.8330960, 1.7357707, 1.1936760], atol=1e-6)
    assert np.allclose(c["reg"], [-1.2863572, -1.2134285, -1.3106679, -1.8520957], atol=1e-6)
    assert np.allclose(CASES[1]["raw"], [0.0431132, 0.5801919, 0.1068089, 0.1194915], atol=1e-6)


@pytest.fixture(scope="module")
def eng():
    from softtreemax_b200.engine import Engine
    e = Engine("synth_breakout")
    yield e
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(CASES)))
def test_kernel_matches_reference_vectors(eng, i):
    from softtreemax_b200 import ops
    c = CASES[i]
    head, rew, d0 = case_inputs(c)
    dev = eng.device
    grp = case_group(c)
    if c["kind"] == "width":
        d0 = None
    logits, raw, _ = ops.softtreemax_logits_full(eng, head[None].to(dev), rew[None].to(dev), c["gamma"] ** c["d"], c["beta"],
                                                c["alpha"], c["eps"], c["cumulative"],
                                                d0.to(dev) if d0 is not None else None,
                                                group=grp.to(dev) if grp is not None else None)
    assert np.allclose(raw[0].cpu().numpy(), want_raw(c), atol=ATOL)
    assert np.allclose(logits[0].cpu().numpy(), c["reg"], atol=ATOL)
    if c["kind"] == "width":
        h = head[None].to(dev).requires_grad_()
        beta = torch.tensor(c["beta"], device=dev, requires_grad=True)
        y = ops.softtreemax_logits(eng, h, rew[None].to(dev), c["gamma"] ** c["d"], beta, 1.0, c["eps"], c["cumulative"],
                                   None, grp.to(dev))
        (y * torch.tensor(c["dy"], device=dev).reshape(1, -1)).sum().backward()
        assert np.allclose(h.grad.flatten().cpu().numpy(), c["dhead"], rtol=1e-4, atol=1e-6)
        if np.isnan(c["dbeta"]):
            # C-mode with an action that has no surviving leaf: the reference divides 0 by a zero count
            # before masking the entry to -inf (actor_critic_ts.py:81,86), so ITS autograd returns
            # dbeta = NaN.  The kernel never forms 0/0 and returns the finite gradient of the live groups.
            assert np.isfinite(float(beta.grad))
        else:
            assert np.isclose(float(beta.grad), c["dbeta"], rtol=1e-4, atol=1e-6)
    if c["kind"] == "random":
        h = head[None].to(dev).requires_grad_()
        dd = d0.to(dev).requires_grad_()
        beta = torch.tensor(c["beta"], device=dev, requires_grad=True)
        alpha = torch.tensor(c["alpha"], device=dev, requires_grad=True)
        y = ops.softtreemax_logits(eng, h, rew[None].to(dev), c["gamma"] ** c["d"], beta, alpha, c["eps"], c["cumulative"], dd)
        (y * torch.tensor(c["dy"], device=dev).reshape(1, -1)).sum().backward()
        assert np.allclose(h.grad.flatten().cpu().numpy(), c["dhead"], rtol=1e-4, atol=1e-6)
        assert np.allclose(dd.grad.flatten().cpu().numpy(), c["ddepth0"], rtol=1e-4, atol=1e-6)
        assert np.isclose(float(beta.grad), c["dbeta"], rtol=1e-4, atol=1e-6)
        assert np.isclose(float(alpha.grad), c["dalpha"], rtol=1e-4, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("A,d,R,cum", [(4, 2, 37, False), (4, 3, 5, True), (6, 3, 3, False), (9, 2, 11, False), (18, 1, 4, True)])
def test_kernel_matches_restatement_batched(eng, A, d, R, cum):
    from oracle import policy_ref
    from softtreemax_b200 import ops
    g = torch.Generator().manual_seed(A * 100 + d)
    L = A ** d
    head = 4 * torch.randn(R, L, A, generator=g)
    rew = torch.randint(-3, 4, (R, L), generator=g).float()
    vals = torch.randn(R, L, generator=g)
    d0 = torch.randn(R, A, generator=g)
    want = policy_ref.softtreemax_logits(head, rew, A, 0.99 ** d, 1.3, 0.7, 0.001, cum, d0, vals)
    dev = eng.device
    got = ops.softtreemax_logits_full(eng, head.to(dev), rew.to(dev), 0.99 ** d, 1.3, 0.7, 0.001, cum, d0.to(dev), vals.to(dev))
    for w, gt in zip(want, got):
        assert torch.allclose(gt.cpu(), w, atol=ATOL, rtol=1e-5)
    # extreme magnitudes stay finite (stable logsumexp)
    big = ops.softtreemax_logits_full(eng, (head * 1e4).to(dev), rew.to(dev), 1.0, 1.0, 1.0, 0.001, False)
    assert torch.isfinite(big[0]).all()
