"""Training step: the differentiable teacher-forced log-likelihood of
rl4co_slap_b200/training.py against the oracle (CPU, values and autograd gradients) and against the
rollout kernels' own log-likelihood (GPU), plus one REINFORCE step end to end."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from rl4co_slap_b200.training import reinforce_step, replay_masks, teacher_forced_log_likelihood
from tests.helpers import env_of, fixture, gen_td, load_policy, oracle_reset, oracle_weights, t

RTOL = 1e-5


def _cpu_td(name, fx):
    if name == "cvrp":
        return {"locs": t(fx["reset.locs"]), "demand": t(fx["gen.demand"]),
                "vehicle_capacity": t(fx["reset.vehicle_capacity"])}
    if name == "op":
        return {"locs": t(fx["reset.locs"]), "prize": t(fx["reset.prize"]), "max_length": t(fx["reset.max_length"])}
    if name in ("pctsp", "spctsp"):
        return {"locs": t(fx["reset.locs"]), "real_prize": t(fx["reset.real_prize"]),
                "prize_required": t(fx["reset.prize_required"])}
    return {"locs": t(fx["gen.locs"])}


@pytest.mark.parametrize("case", ("pctsp20", "pctsp100", "spctsp50"))
def test_teacher_forced_ll_pctsp_cpu(case):
    """(s)pctsp: masks, collected prize and per-step log-probs of the reference's greedy tours and of random tours"""
    fx, name = fixture(case), env_of(case)
    pol = load_policy(name, "cpu")
    acts = t(fx["replay.actions"])
    mask, valid, total = replay_masks(name, _cpu_td(name, fx), acts)
    assert np.array_equal(mask[:, 0].numpy(), fx["reset.action_mask"])
    for s in range(acts.shape[1] - 1):
        assert np.array_equal(mask[:, s + 1].numpy(), fx[f"step{s}.action_mask"]), s
        assert np.array_equal(total[:, s + 1].numpy(), fx[f"step{s}.cur_total_prize"]), s
    with torch.no_grad():
        lp = teacher_forced_log_likelihood(pol, _cpu_td(name, fx), name, t(fx["hidden"]), acts, return_sum=False)
        lg = teacher_forced_log_likelihood(pol, _cpu_td(name, fx), name, t(fx["hidden"]), t(fx["greedy.actions"]),
                                           return_sum=False)
    np.testing.assert_allclose(lp.numpy(), fx["eval.logprobs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(lg.numpy(), fx["greedy.logprobs"], rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("case", ("op20", "op50", "op100"))
def test_teacher_forced_ll_op_cpu(case):
    """op: per-step log-probs of the replay's random tours (and the masks behind them) equal the reference's"""
    fx = fixture(case)
    pol = load_policy("op", "cpu")
    acts = t(fx["replay.actions"])
    mask, valid, length = replay_masks("op", _cpu_td("op", fx), acts)
    assert np.array_equal(mask[:, 0].numpy(), fx["reset.action_mask"])
    for s in range(acts.shape[1] - 1):
        assert np.array_equal(mask[:, s + 1].numpy(), fx[f"step{s}.action_mask"]), s
        assert np.array_equal(length[:, s + 1].numpy(), fx[f"step{s}.tour_length"]), s
    with torch.no_grad():
        lp = teacher_forced_log_likelihood(pol, _cpu_td("op", fx), "op", t(fx["hidden"]), acts, return_sum=False)
    np.testing.assert_allclose(lp.numpy(), fx["eval.logprobs"], rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("case", ("tsp20", "tsp50", "cvrp20", "cvrp50", "slap50"))
def test_teacher_forced_ll_matches_oracle_cpu(case):
    """values: the batched all-steps pass equals the oracle's step-by-step evaluate of the reference's tours"""
    fx, name = fixture(case), env_of(case)
    pol = load_policy(name, "cpu")
    h, acts = t(fx["hidden"]), t(fx["greedy.actions"])
    with torch.no_grad():
        lp = teacher_forced_log_likelihood(pol, _cpu_td(name, fx), name, h, acts, return_sum=False)
    np.testing.assert_allclose(lp.numpy(), fx["greedy.logprobs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(lp.sum(1).numpy(), fx["greedy.ll"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("case", ("tsp20", "cvrp20"))
def test_teacher_forced_ll_multistart_cpu(case):
    fx, name = fixture(case), env_of(case)
    pol = load_policy(name, "cpu")
    S = int(fx["ms_greedy.num_starts"])
    with torch.no_grad():
        ll = teacher_forced_log_likelihood(pol, _cpu_td(name, fx), name, t(fx["hidden"]), t(fx["ms_greedy.actions"]),
                                           num_starts=S)
    np.testing.assert_allclose(ll.numpy(), fx["ms_greedy.ll"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("case", ("tsp20", "cvrp20"))
def test_replayed_masks_match_reference_trace(case):
    """the masks fed to the decoder are the reference's per-step action masks bit for bit"""
    fx, name = fixture(case), env_of(case)
    acts = t(fx["actions"]) if "actions" in fx.files else t(fx["greedy.actions"])
    mask, valid, used = replay_masks(name, _cpu_td(name, fx), acts)
    assert np.array_equal(mask[:, 0].numpy(), fx["reset.action_mask"].astype(bool))
    for s in range(min(acts.shape[1] - 1, 6)):
        if f"step{s}.action_mask" in fx.files and "actions" in fx.files:
            assert np.array_equal(mask[:, s + 1].numpy(), fx[f"step{s}.action_mask"].astype(bool)), s


@pytest.mark.parametrize("case", ("tsp20", "cvrp20", "slap50"))
def test_gradients_match_oracle_autograd_cpu(case):
    """d ll / d (decoder weights, h) of the batched pass equals autograd through the oracle's reference-order loop"""
    fx, name = fixture(case), env_of(case)
    pol = load_policy(name, "cpu").double()
    h = t(fx["hidden"]).double().requires_grad_(True)
    acts = t(fx["greedy.actions"])
    # decoder in float64, environment state (capacities, masks) in float32 like the reference's
    ll = teacher_forced_log_likelihood(pol, _cpu_td(name, fx), name, h, acts)
    ll.sum().backward()
    W = {k: v.double().requires_grad_(True) for k, v in oracle_weights(name).items()}
    h2 = t(fx["hidden"]).double().requires_grad_(True)
    ref = port.am_rollout(name, oracle_reset(name, fx), h2, W, actions=acts, check_solution=False)["log_likelihood"]
    ref.sum().backward()
    torch.testing.assert_close(ll.detach(), ref.detach(), rtol=1e-9, atol=1e-9)
    torch.testing.assert_close(h.grad, h2.grad, rtol=1e-7, atol=1e-9)
    d = pol.decoder
    torch.testing.assert_close(d.project_node_embeddings.weight.grad, W["Wnode"].grad, rtol=1e-7, atol=1e-9)
    torch.testing.assert_close(d.pointer.project_out.weight.grad, W["Wout"].grad, rtol=1e-7, atol=1e-9)
    torch.testing.assert_close(d.project_fixed_context.weight.grad, W["Wfixed"].grad, rtol=1e-7, atol=1e-9)
    if name != "slap":
        torch.testing.assert_close(d.context_embedding.project_context.weight.grad, W["Wctx"].grad, rtol=1e-7, atol=1e-9)
    if name == "tsp":
        torch.testing.assert_close(d.context_embedding.W_placeholder.grad, W["Wph"].grad, rtol=1e-7, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("name,decode_type,S", (("tsp", "sampling", 0), ("cvrp", "sampling", 0),
                                                ("cvrp", "multistart_sampling", 7), ("tsp", "multistart_sampling", 5)))
def test_differentiable_ll_matches_kernel_ll(name, decode_type, S, dev):
    """on sampled trajectories the differentiable pass reproduces the kernel's log-likelihood"""
    torch.manual_seed(11)
    pol, env = load_policy(name, dev), R.get_env(name, device=dev, generator_params=dict(num_loc=20))
    td = env.reset(batch_size=[48])
    with torch.no_grad():
        out = pol(td, env, decode_type=decode_type, num_starts=S or None, return_actions=True, return_hidden=True)
        ll = teacher_forced_log_likelihood(pol, td, name, out["hidden"].node_embeddings, out["actions"], num_starts=S)
    torch.testing.assert_close(ll, out["log_likelihood"], rtol=RTOL, atol=2e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("name,decode_type,baseline", (("tsp", "sampling", "mean"), ("cvrp", "multistart_sampling", "shared")))
def test_reinforce_step_trains(name, decode_type, baseline, dev):
    """one REINFORCE step on a fixed batch: finite loss, the differentiable ll equals the kernel's, gradients reach
    encoder and decoder, and (plain SGD, small step) the surrogate objective mean(advantage * ll) of the SAME sampled
    trajectories goes up - the first-order guarantee of a correct policy gradient"""
    torch.manual_seed(3)
    pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=2).to(dev)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=20))
    td = env.reset(batch_size=[64])
    S = 6 if baseline == "shared" else 0
    opt = torch.optim.SGD(pol.parameters(), lr=1e-3)
    before = [p.detach().clone() for p in pol.parameters()]
    with torch.no_grad():
        probe = pol(td, env, decode_type=decode_type, num_starts=S or None)
    bl = "shared" if baseline == "shared" else probe["reward"].mean()
    first = reinforce_step(pol, env, td, opt, baseline=bl, decode_type=decode_type, num_starts=S or None, max_grad_norm=None)
    assert torch.isfinite(first["loss"])
    torch.testing.assert_close(first["log_likelihood"], first["kernel_log_likelihood"], rtol=1e-4, atol=1e-4)
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in pol.parameters() if p.requires_grad)
    changed = sum(int(not torch.equal(a, b.detach())) for a, b in zip(before, pol.parameters()))
    assert changed >= len(before) - 2, "the optimiser step did not reach (almost) every parameter"
    reward = first["reward"]
    adv = reward - (reward.reshape(max(S, 1), -1).mean(0, keepdim=True).expand(max(S, 1), -1).reshape(-1) if S else bl)
    with torch.no_grad():
        h, _ = pol.encoder(td)
        ll_after = teacher_forced_log_likelihood(pol, td, name, h.float(), first["actions"], num_starts=S)
    j0, j1 = (adv * first["log_likelihood"]).mean().item(), (adv * ll_after).mean().item()
    assert j1 > j0, f"surrogate objective did not increase: {j0} -> {j1}"


@pytest.mark.gpu
def test_small_baselines_follow_reference_formulas(dev):
    """No / Exponential / Mean / Warmup baselines (baselines.py:49-135) and the registry; the exponential baseline's
    batch mean comes from the fp64 partial-sum kernel"""
    torch.manual_seed(0)
    r1, r2 = torch.randn(4096, device=dev) - 7, torch.randn(4096, device=dev) - 6
    assert R.get_reinforce_baseline("no").eval(None, r1) == (0, 0)
    eb = R.get_reinforce_baseline("exponential", beta=0.8)
    v1, l1 = eb.eval(None, r1)
    v2, _ = eb.eval(None, r2)
    torch.testing.assert_close(v1, r1.double().mean().float(), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(v2, (0.8 * r1.double().mean() + 0.2 * r2.double().mean()).float(), rtol=1e-6, atol=1e-6)
    assert l1 == 0 and v2.dtype == torch.float32 and not v2.requires_grad
    mb = R.get_reinforce_baseline("mean")
    mb.eval(None, r1)
    torch.testing.assert_close(mb.eval(None, r2)[0], r2.double().mean().float(), rtol=1e-6, atol=1e-6)
    wb = R.WarmupBaseline(R.get_reinforce_baseline("mean"), n_epochs=2, warmup_exp_beta=0.5)
    a0 = wb.eval(None, r1)[0]                                   # alpha 0: the warm-up exponential baseline alone
    torch.testing.assert_close(a0, r1.double().mean().float(), rtol=1e-6, atol=1e-6)
    wb.epoch_callback(epoch=0)
    assert wb.alpha == 0.5
    mix = wb.eval(None, r2)[0]
    want = 0.5 * r2.double().mean() + 0.5 * (0.5 * r1.double().mean() + 0.5 * r2.double().mean())
    torch.testing.assert_close(mix, want.float(), rtol=1e-6, atol=1e-6)
    wb.epoch_callback(epoch=1)
    assert wb.alpha == 1.0
    assert isinstance(R.get_reinforce_baseline("rollout"), R.WarmupBaseline)
    with pytest.raises(ValueError, match="Unknown baseline"):
        R.get_reinforce_baseline("critic_nope")
    # a baseline object drives reinforce_step exactly like the reference's calculate_loss (reinforce.py:95-106)
    pol = R.AttentionModelPolicy(env_name="tsp", num_encoder_layers=1).to(dev)
    env = R.TSPEnv(generator_params=dict(num_loc=10), device=dev)
    td = env.reset(batch_size=[32])
    res = reinforce_step(pol, env, td, optimizer=None, baseline=R.get_reinforce_baseline("exponential"))
    adv = res["reward"] - res["reward"].double().mean().float()
    torch.testing.assert_close(res["loss"], -(adv * res["log_likelihood"]).mean(), rtol=1e-5, atol=1e-6)
    # SharedBaseline objects (from the class, the registry, and inside a warm-up baseline) through reinforce_step
    S = 5
    for bl in (R.SharedBaseline(), R.get_reinforce_baseline("shared"),
               R.get_reinforce_baseline("warmup", baseline="shared", n_epochs=1)):
        if isinstance(bl, R.WarmupBaseline):
            bl.epoch_callback(epoch=0)      # alpha = 1: the shared baseline alone
        res = reinforce_step(pol, env, td, optimizer=None, baseline=bl, decode_type="multistart_sampling", num_starts=S)
        r2 = R.unbatchify(res["reward"], S)
        adv = (r2 - r2.mean(1, keepdim=True)).t().reshape(-1)
        torch.testing.assert_close(res["loss"], -(adv * res["log_likelihood"]).mean(), rtol=1e-5, atol=1e-6)
