"""GPU parity of the prize-collecting TSP and its stochastic variant (sibling routing environments, SURVEY.md 8f
rank 4): env kernels and the register-file rollout kernel vs the fixtures of the unmodified reference and the oracle
port.  (tests/test_gpu_envs.py runs the step-by-step and fused replays of the fixtures bit-exact.)
Tours / masks bit-exact; rewards and log-likelihoods within 1e-5 relative."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import PCTSP_CASES, env_of, fixture, gen_td, load_policy, t

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _setup(case, dev):
    fx, name = fixture(case), env_of(case)
    pol, env = load_policy(name, dev), R.get_env(name, device=dev)
    return fx, pol, env, env.reset(gen_td(name, fx, dev))


@pytest.mark.parametrize("case", PCTSP_CASES)
def test_greedy_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    out = pol(td, env, decode_type="greedy", return_actions=True, hidden=t(fx["hidden"], dev),
              return_sum_log_likelihood=False)
    assert np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"]), "tours differ from the reference"
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["greedy.reward"], rtol=RTOL, atol=2e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["greedy.logprobs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out["log_likelihood"].sum(1).cpu().numpy(), fx["greedy.ll"], rtol=RTOL, atol=2e-5)


@pytest.mark.parametrize("case", PCTSP_CASES)
def test_evaluate_random_tours_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    acts = t(fx["replay.actions"], dev)
    out = pol(td, env, actions=acts, hidden=t(fx["hidden"], dev), return_actions=True,
              return_sum_log_likelihood=False)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["eval.logprobs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out["log_likelihood"].sum(1).cpu().numpy(), fx["eval.ll"], rtol=RTOL, atol=3e-5)
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["eval.reward"], rtol=RTOL, atol=2e-6)


@pytest.mark.parametrize("name,size,B", [("pctsp", 100, 48), ("spctsp", 127, 9), ("pctsp", 33, 40), ("spctsp", 5, 16)])
def test_full_policy_vs_oracle(name, size, B, dev):
    """random instances, fresh random weights, the PyTorch encoder on the GPU: greedy tours against the oracle fed
    with the same embeddings; the final state of the fused replay against the oracle's"""
    torch.manual_seed(200 + size)
    env = R.get_env(name, generator_params=dict(num_loc=size), device=dev)
    pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=2).to(dev).eval()
    td0 = env.generator(B)
    td = env.reset(td0)
    out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True, return_sum_log_likelihood=False)
    h = out["hidden"].node_embeddings.cpu()
    sd = {k: v.detach().cpu() for k, v in pol.state_dict().items()}
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"], "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"], "Wctx": sd["decoder.context_embedding.project_context.weight"]}
    c = td0.cpu()
    o0 = port.pctsp_reset(c["depot"], c["locs"], c["penalty"], c["deterministic_prize"], c["stochastic_prize"],
                          stochastic=name == "spctsp")
    assert torch.equal(td["action_mask"].cpu(), o0["action_mask"])
    ref = port.am_rollout(name, o0, h, W, "greedy")
    assert torch.equal(out["actions"].cpu(), ref["actions"])
    torch.testing.assert_close(out["log_likelihood"].cpu(), ref["logprobs"], rtol=1e-4, atol=3e-5)
    torch.testing.assert_close(out["reward"].cpu(), ref["reward"], rtol=RTOL, atol=2e-6)
    final, reward = env.replay(td, out["actions"])
    torch.testing.assert_close(reward.cpu(), ref["reward"], rtol=RTOL, atol=2e-6)
    for k in ("action_mask", "visited", "cur_total_prize", "current_node", "i", "done"):
        assert torch.equal(final[k].cpu().view(ref["td"][k].shape), ref["td"][k]), k
    torch.testing.assert_close(final["cur_total_penalty"].cpu(), ref["td"]["cur_total_penalty"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("name", ("pctsp", "spctsp"))
def test_sampling_is_feasible_and_consistent(name, dev):
    torch.manual_seed(11)
    env = R.get_env(name, generator_params=dict(num_loc=50), device=dev)
    pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(512))
    out = pol(td, env, decode_type="sampling", return_actions=True, return_hidden=True)
    env.check_solution_validity(td, out["actions"])
    c = {k: v.cpu() for k, v in td.items()}
    port.pctsp_check(c, out["actions"].cpu())
    ev = pol(td, env, actions=out["actions"], hidden=out["hidden"].node_embeddings)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=RTOL, atol=3e-5)
    torch.testing.assert_close(ev["reward"], out["reward"], rtol=RTOL, atol=2e-6)
    torch.testing.assert_close(out["reward"].cpu(), port.pctsp_reward(c, out["actions"].cpu()), rtol=RTOL, atol=2e-6)
    out2 = pol(td, env, decode_type="sampling", return_actions=True)
    T = min(out["actions"].shape[1], out2["actions"].shape[1])
    assert (out["actions"][:, :T] != out2["actions"][:, :T]).any()


def test_invalid_tours_raise_reference_messages(dev):
    fx, pol, env, td = _setup("pctsp20", dev)
    acts = t(fx["replay.actions"], dev).clone()
    bad = acts.clone()
    bad[0, 1] = bad[0, 0]
    with pytest.raises(AssertionError, match="Duplicates"):
        env.check_solution_validity(td, bad)
    short = torch.zeros_like(acts)
    short[:, 0] = 1                                      # one customer, then the depot: far too little prize
    with pytest.raises(AssertionError, match="Total prize does not satisfy min total prize"):
        env.check_solution_validity(td, short)


@pytest.mark.parametrize("case", PCTSP_CASES)
def test_multistart_greedy_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    S = int(fx["ms_greedy.num_starts"])
    out = pol(td, env, decode_type="multistart_greedy", num_starts=S, return_actions=True, hidden=t(fx["hidden"], dev))
    assert np.array_equal(out["actions"].cpu().numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["ms_greedy.reward"], rtol=RTOL, atol=2e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["ms_greedy.ll"], rtol=RTOL, atol=3e-5)


def test_pomo_sampling_default_starts(dev):
    """multistart sampling with the default number of starts (one per customer): valid tours, ll = evaluate ll, and
    the shared-baseline REINFORCE step's differentiable ll = the kernel's"""
    torch.manual_seed(21)
    env = R.PCTSPEnv(generator_params=dict(num_loc=20), device=dev)
    pol = R.AttentionModelPolicy(env_name="pctsp", num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(32))
    res = R.reinforce_step(pol, env, td, optimizer=None, decode_type="multistart_sampling", baseline="shared")
    assert res["actions"].shape[0] == 32 * 20
    first = res["actions"][:, 0].reshape(20, 32)
    assert torch.equal(first, (torch.arange(20, device=dev) + 1)[:, None].expand(20, 32))
    torch.testing.assert_close(res["log_likelihood"], res["kernel_log_likelihood"], rtol=1e-4, atol=5e-5)
    from rl4co_slap_b200.ops import batchify
    env.check_solution_validity(batchify(td, 20), res["actions"])


def test_training_step_log_likelihood(dev):
    torch.manual_seed(5)
    env = R.PCTSPEnv(generator_params=dict(num_loc=20), device=dev)
    pol = R.AttentionModelPolicy(env_name="pctsp", num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(64))
    res = R.reinforce_step(pol, env, td, optimizer=None, decode_type="sampling")
    torch.testing.assert_close(res["log_likelihood"], res["kernel_log_likelihood"], rtol=1e-4, atol=5e-5)
