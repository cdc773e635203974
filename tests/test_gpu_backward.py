"""The backward kernel of the teacher-forced log-likelihood (csrc/am_backward.cu) against autograd
through the batched PyTorch pass (training.teacher_forced_log_likelihood, itself pinned on the oracle's autograd in
tests/test_training.py): log-likelihoods and the gradients with respect to the encoder output and EVERY decoder weight,
tsp and cvrp, single start and POMO multistart, with and without graph context, up to N = 100."""
import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200.training import (kernel_backward_supported, kernel_log_likelihood, reinforce_step,
                                      teacher_forced_log_likelihood)
from tests.helpers import load_policy

pytestmark = pytest.mark.gpu


def _grads(fn, pol, h, weights):
    h = h.detach().clone().requires_grad_(True)
    params = [p for p in pol.decoder.parameters() if p.requires_grad]
    for p in params:
        p.grad = None
    ll = fn(h)
    (weights * ll).sum().backward()
    out = {"h": h.grad.detach().clone()}
    for name, p in pol.decoder.named_parameters():
        if p.requires_grad:
            out[name] = None if p.grad is None else p.grad.detach().clone()
    return ll.detach(), out


@pytest.mark.parametrize("name,num_loc,B,S,variant", (
    ("tsp", 20, 24, 0, None), ("tsp", 20, 9, 6, None), ("tsp", 100, 5, 0, None), ("tsp", 50, 3, 50, None),
    ("cvrp", 20, 24, 0, None), ("cvrp", 20, 9, 6, None), ("cvrp", 100, 4, 0, None), ("cvrp", 20, 10, 20, "cvrp_pomo")))
def test_backward_kernel_matches_autograd(name, num_loc, B, S, variant, dev):
    torch.manual_seed(500 + num_loc + S)
    pol = load_policy(name, dev, variant)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=num_loc))
    td = env.reset(env.generator(B).to(dev))
    with torch.no_grad():
        h0 = pol.encoder(td)[0].float()
        kw = dict(decode_type="multistart_sampling", num_starts=S) if S else dict(decode_type="sampling")
        out = pol(td, env, return_actions=True, hidden=h0, **kw)
    actions = out["actions"]
    assert kernel_backward_supported(pol, name, h0)
    w = torch.randn(actions.shape[0], device=dev)          # dL / d ll of every rollout (the REINFORCE advantages)
    ll_ref, g_ref = _grads(lambda h: teacher_forced_log_likelihood(pol, td, name, h, actions, num_starts=S), pol, h0, w)
    ll_ker, g_ker = _grads(lambda h: kernel_log_likelihood(pol, td, name, h, actions, num_starts=S), pol, h0, w)
    torch.testing.assert_close(ll_ker, ll_ref, rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(ll_ker, out["log_likelihood"], rtol=1e-5, atol=2e-5)
    assert set(g_ker) == set(g_ref)
    for key, want in g_ref.items():
        got = g_ker[key]
        if want is None or float(want.abs().max()) == 0.0:      # a weight the pass does not touch (e.g. no graph context)
            assert got is None or float(got.abs().max()) < 1e-6, key
            continue
        assert got is not None, key
        scale = float(want.abs().max())
        err = float((got - want).abs().max())
        assert err <= 2e-4 * scale, f"{key}: max |grad diff| {err:.3e} against a gradient of size {scale:.3e}"


def test_reinforce_step_uses_the_backward_kernel(dev):
    """reinforce_step with backward='kernel' and backward='torch' on the same sampled tours: same loss, same gradients"""
    for name, dt, S in (("tsp", "sampling", None), ("cvrp", "multistart_sampling", 5)):
        torch.manual_seed(3)
        pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=2).to(dev)
        env = R.get_env(name, device=dev, generator_params=dict(num_loc=20))
        td = env.reset(batch_size=[32])
        res = {}
        for mode in ("kernel", "torch"):
            torch.manual_seed(11)               # same Philox seed -> same tours
            for p in pol.parameters():
                p.grad = None
            opt = torch.optim.SGD(pol.parameters(), lr=0.0)
            out = reinforce_step(pol, env, td, opt, baseline="shared" if S else None, decode_type=dt, num_starts=S,
                                 max_grad_norm=None, backward=mode)
            res[mode] = (out, {n: p.grad.detach().clone() for n, p in pol.named_parameters() if p.grad is not None})
        assert torch.equal(res["kernel"][0]["actions"], res["torch"][0]["actions"])
        torch.testing.assert_close(res["kernel"][0]["loss"], res["torch"][0]["loss"], rtol=1e-5, atol=1e-6)
        gk, gt = res["kernel"][1], res["torch"][1]
        assert set(gk) == set(gt)
        # (a bias that feeds batch norm has a mathematically zero gradient: both paths hold rounding noise there, so the
        # yardstick is the larger of the parameter's own gradient and a small fraction of the largest gradient overall)
        top = max(float(v.abs().max()) for v in gt.values())
        for key in gt:
            scale = max(float(gt[key].abs().max()), 1e-2 * top)
            assert float((gk[key] - gt[key]).abs().max()) <= 3e-4 * scale, key
