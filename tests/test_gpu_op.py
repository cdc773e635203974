"""GPU parity of the orienteering problem (OP), the sibling routing environment:
env kernels and the register-file rollout kernel vs the fixtures of the unmodified reference and the oracle
port.  (tests/test_gpu_envs.py runs the step-by-step and fused replays of the op fixtures bit-exact.)
Tours / masks bit-exact; rewards and log-likelihoods within 1e-5 relative."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import OP_CASES, fixture, gen_td, load_policy, oracle_reset, oracle_weights, t

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _setup(case, dev):
    fx = fixture(case)
    pol, env = load_policy("op", dev), R.OPEnv(device=dev)
    return fx, pol, env, env.reset(gen_td("op", fx, dev))


@pytest.mark.parametrize("case", OP_CASES)
def test_greedy_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    out = pol(td, env, decode_type="greedy", return_actions=True, hidden=t(fx["hidden"], dev),
              return_sum_log_likelihood=False)
    assert np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"]), "tours differ from the reference"
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["greedy.reward"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["greedy.logprobs"], rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("case", OP_CASES)
def test_evaluate_random_tours_matches_reference_fixture(case, dev):
    """the reference policy's per-step log-probs of the replay's random tours (they run into the length budget,
    so the masks that feed the glimpse and the log-softmax are exercised)"""
    fx, pol, env, td = _setup(case, dev)
    acts = t(fx["replay.actions"], dev)
    out = pol(td, env, actions=acts, hidden=t(fx["hidden"], dev), return_actions=True,
              return_sum_log_likelihood=False)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["eval.logprobs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out["log_likelihood"].sum(1).cpu().numpy(), fx["eval.ll"], rtol=RTOL, atol=2e-5)
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["eval.reward"], rtol=RTOL, atol=1e-6)


def _nearest_neighbour_tours(o0):
    """long feasible tours from the oracle env: always the nearest feasible customer, depot when none is left"""
    td = dict(o0)
    acts = []
    while not td["done"].all():
        cur = port.gather_by_index(td["locs"], td["current_node"])[:, None, :]
        d = (td["locs"] - cur).norm(p=2, dim=-1).masked_fill(~td["action_mask"], float("inf"))
        d[:, 0] = 1e9                                    # depot only when nothing else is feasible
        a = d.argmin(-1)
        td["action"] = a
        td = port.op_step(td)
        acts.append(a)
    return torch.stack(acts, 1)


@pytest.mark.parametrize("size,B", [(100, 48), (127, 9), (33, 40), (5, 16)])
def test_full_policy_vs_oracle(size, B, dev):
    """random instances, fresh random weights, the PyTorch encoder on the GPU: greedy tours and nearest-neighbour
    tours (the longest the budget allows) against the oracle fed with the same embeddings"""
    torch.manual_seed(100 + size)
    env = R.OPEnv(generator_params=dict(num_loc=size), device=dev)
    pol = R.AttentionModelPolicy(env_name="op", num_encoder_layers=2).to(dev).eval()
    td0 = env.generator(B)
    td = env.reset(td0)
    out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True, return_sum_log_likelihood=False)
    h = out["hidden"].node_embeddings.cpu()
    sd = {k: v.detach().cpu() for k, v in pol.state_dict().items()}
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"], "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"], "Wctx": sd["decoder.context_embedding.project_context.weight"]}
    c = td0.cpu()
    o0 = port.op_reset(c["depot"], c["locs"], c["prize"], c["max_length"])
    assert torch.equal(td["max_length"].cpu(), o0["max_length"]), "reset budgets differ from the CPU reference arithmetic"
    assert torch.equal(td["action_mask"].cpu(), o0["action_mask"])
    ref = port.am_rollout("op", o0, h, W, "greedy")
    assert torch.equal(out["actions"].cpu(), ref["actions"])
    torch.testing.assert_close(out["log_likelihood"].cpu(), ref["logprobs"], rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(out["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-6)
    nn_acts = _nearest_neighbour_tours(o0)
    assert nn_acts.shape[1] > min(size, 8)              # these are long tours
    ref = port.am_rollout("op", o0, h, W, actions=nn_acts)
    got = pol(td, env, actions=nn_acts.to(dev), hidden=h.to(dev), return_sum_log_likelihood=False)
    torch.testing.assert_close(got["log_likelihood"].cpu(), ref["logprobs"], rtol=1e-4, atol=3e-5)
    torch.testing.assert_close(got["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-6)
    final, reward = env.replay(td, nn_acts.to(dev))
    torch.testing.assert_close(reward.cpu(), ref["reward"], rtol=RTOL, atol=1e-6)
    for k in ("action_mask", "visited", "tour_length", "current_total_prize", "current_node", "i", "done"):
        assert torch.equal(final[k].cpu().view(ref["td"][k].shape), ref["td"][k]), k


def test_sampling_is_feasible_and_consistent(dev):
    """sampled tours are valid (the reference's own validity check), their log-likelihood equals the evaluate
    path's on the same actions, and repeated sampling visits different tours"""
    torch.manual_seed(7)
    env = R.OPEnv(generator_params=dict(num_loc=50), device=dev)
    pol = R.AttentionModelPolicy(env_name="op", num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(512))
    out = pol(td, env, decode_type="sampling", return_actions=True, return_hidden=True)
    env.check_solution_validity(td, out["actions"])
    c = {k: v.cpu() for k, v in td.items()}
    port.op_check(c["locs"], c["max_length"], out["actions"].cpu())
    ev = pol(td, env, actions=out["actions"], hidden=out["hidden"].node_embeddings)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=RTOL, atol=2e-5)
    torch.testing.assert_close(ev["reward"], out["reward"], rtol=RTOL, atol=1e-6)
    torch.testing.assert_close(out["reward"].cpu(), port.op_reward(c["prize"], out["actions"].cpu()), rtol=RTOL, atol=1e-6)
    out2 = pol(td, env, decode_type="sampling", return_actions=True)
    T = min(out["actions"].shape[1], out2["actions"].shape[1])
    assert (out["actions"][:, :T] != out2["actions"][:, :T]).any()


def test_invalid_tours_raise_reference_messages(dev):
    fx, pol, env, td = _setup("op20", dev)
    acts = t(fx["replay.actions"], dev).clone()
    bad = acts.clone()
    row = int((bad[:, 1] != 0).nonzero()[0])
    bad[row, 0] = bad[row, 1]
    with pytest.raises(AssertionError, match="Duplicates"):
        env.check_solution_validity(td, bad)
    # every customer in index order is far beyond the length budget
    N = td["locs"].shape[1]
    long = torch.cat((torch.arange(1, N, device=dev), torch.zeros(1, dtype=torch.int64, device=dev)))[None].repeat(acts.shape[0], 1)
    with pytest.raises(AssertionError, match="Max length exceeded"):
        env.check_solution_validity(td, long)
    with pytest.raises(AssertionError, match="Max length exceeded"):     # get_reward's validity check, as the reference
        pol(td, env, actions=long, hidden=t(fx["hidden"], dev))


def test_multistart_start_nodes_are_resampled_like_the_reference(dev):
    """op10 has instances with fewer feasible nodes than starts: the reference draws the start nodes from the feasible
    ones with torch.multinomial (utils/ops.py:166-176).  The policy does the same draw (same call, torch's generator) and
    hands the nodes to the kernel; the oracle given the SAME nodes must produce the same tours."""
    fx, pol, env, td = _setup("op10", dev)
    S, B = 10, td["locs"].shape[0]
    feasible = td["action_mask"][..., 1:]
    assert bool((feasible.sum(-1) < S).any()), "the fixture is meant to have instances with too few feasible starts"
    torch.manual_seed(5)
    out = pol(td, env, decode_type="multistart_greedy", num_starts=S, hidden=t(fx["hidden"], dev), return_actions=True)
    torch.manual_seed(5)        # the draw the policy made
    sel = (torch.multinomial(feasible.float(), S, replacement=True) + 1).t().reshape(-1)
    assert torch.equal(out["actions"][:, 0], sel)
    assert feasible[torch.arange(B, device=dev).repeat(S), sel - 1].all(), "a start node is infeasible"
    ref = port.am_rollout("op", oracle_reset("op", fx), t(fx["hidden"]), oracle_weights("op"), "multistart_greedy",
                          num_starts=S, start_nodes=sel.cpu())
    got = out["actions"].cpu()
    Tm = max(got.shape[1], ref["actions"].shape[1])
    pad = lambda x: torch.nn.functional.pad(x, (0, Tm - x.shape[1]))
    assert torch.equal(pad(got), pad(ref["actions"]))
    torch.testing.assert_close(out["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-6)
    torch.testing.assert_close(out["log_likelihood"].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)


def test_select_start_nodes_fn_is_honoured(dev):
    """utils/decoding.py:286-291: a caller's select_start_nodes_fn(td, env, num_starts) replaces the default rule"""
    fx, pol, env, td = _setup("op20", dev)
    S, B = 6, td["locs"].shape[0]
    calls = []

    def backwards(td_, env_, n):
        calls.append(n)
        return (20 - torch.arange(n, device=dev).repeat_interleave(td_["locs"].shape[0]) % 20)   # customers 20, 19, ...

    out = pol(td, env, decode_type="multistart_greedy", num_starts=S, hidden=t(fx["hidden"], dev), return_actions=True,
              select_start_nodes_fn=backwards)
    assert calls == [S]
    assert torch.equal(out["actions"][:, 0].view(S, B), (20 - torch.arange(S, device=dev))[:, None].expand(S, B))
    ref = port.am_rollout("op", oracle_reset("op", fx), t(fx["hidden"]), oracle_weights("op"), "multistart_greedy",
                          num_starts=S, start_nodes=out["actions"][:, 0].cpu(), check_solution=False)
    Tm = max(out["actions"].shape[1], ref["actions"].shape[1])
    pad = lambda x: torch.nn.functional.pad(x, (0, Tm - x.shape[1]))
    assert torch.equal(pad(out["actions"].cpu()), pad(ref["actions"]))
    torch.testing.assert_close(out["log_likelihood"].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)
    with pytest.raises(ValueError, match="rollouts"):
        pol(td, env, decode_type="multistart_greedy", num_starts=S, hidden=t(fx["hidden"], dev),
            select_start_nodes_fn=lambda a, b, n: torch.ones(3, dtype=torch.int64, device=dev))


@pytest.mark.parametrize("case", ("op20", "op50", "op100"))
def test_multistart_greedy_matches_reference_fixture(case, dev):
    """POMO starts (forced first customer s % num_loc + 1, log-prob 0) where the reference does not resample them"""
    fx, pol, env, td = _setup(case, dev)
    S = int(fx["ms_greedy.num_starts"])
    out = pol(td, env, decode_type="multistart_greedy", num_starts=S, return_actions=True, hidden=t(fx["hidden"], dev))
    assert np.array_equal(out["actions"].cpu().numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["ms_greedy.reward"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["ms_greedy.ll"], rtol=RTOL, atol=2e-5)


def test_training_step_log_likelihood(dev):
    """the differentiable teacher-forced pass (training.py) reproduces the kernel's log-likelihood for op"""
    torch.manual_seed(3)
    env = R.OPEnv(generator_params=dict(num_loc=20), device=dev)
    pol = R.AttentionModelPolicy(env_name="op", num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(64))
    res = R.reinforce_step(pol, env, td, optimizer=None, decode_type="sampling")
    torch.testing.assert_close(res["log_likelihood"], res["kernel_log_likelihood"], rtol=1e-4, atol=5e-5)
