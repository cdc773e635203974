"""GPU parity of the split-delivery VRP rollout: the register-file kernel with the per-step
dynamic embedding of the remaining demands vs the fixtures of the unmodified reference and the oracle port.
(tests/test_gpu_envs.py runs the step-by-step and fused replays of the fixtures bit-exact.)"""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import SDVRP_CASES, fixture, gen_td, load_policy, t

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _setup(case, dev):
    fx = fixture(case)
    pol, env = load_policy("sdvrp", dev), R.SDVRPEnv(device=dev)
    return fx, pol, env, env.reset(gen_td("sdvrp", fx, dev))


@pytest.mark.parametrize("case", SDVRP_CASES)
def test_greedy_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    out = pol(td, env, decode_type="greedy", return_actions=True, hidden=t(fx["hidden"], dev),
              return_sum_log_likelihood=False)
    assert np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"]), "tours differ from the reference"
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["greedy.reward"], rtol=RTOL, atol=2e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["greedy.logprobs"], rtol=1e-4, atol=3e-5)


@pytest.mark.parametrize("case", SDVRP_CASES)
def test_evaluate_random_tours_matches_reference_fixture(case, dev):
    fx, pol, env, td = _setup(case, dev)
    acts = t(fx["replay.actions"], dev)
    out = pol(td, env, actions=acts, hidden=t(fx["hidden"], dev), return_actions=True, return_sum_log_likelihood=False)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["eval.logprobs"], rtol=1e-4, atol=3e-5)
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["eval.reward"], rtol=RTOL, atol=2e-6)


@pytest.mark.parametrize("size,B", [(100, 40), (127, 9), (33, 40), (5, 16)])
def test_full_policy_vs_oracle(size, B, dev):
    torch.manual_seed(300 + size)
    env = R.SDVRPEnv(generator_params=dict(num_loc=size), device=dev)
    pol = R.AttentionModelPolicy(env_name="sdvrp", num_encoder_layers=2).to(dev).eval()
    with torch.no_grad():       # a dynamic embedding large enough to matter against the static keys
        pol.decoder.dynamic_embedding.projection.weight.mul_(3.0)
    td0 = env.generator(B)
    td = env.reset(td0)
    out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True, return_sum_log_likelihood=False)
    h = out["hidden"].node_embeddings.cpu()
    sd = {k: v.detach().cpu() for k, v in pol.state_dict().items()}
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"], "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"], "Wctx": sd["decoder.context_embedding.project_context.weight"],
         "Wdyn": sd["decoder.dynamic_embedding.projection.weight"]}
    c = td0.cpu()
    o0 = port.sdvrp_reset(c["depot"], c["locs"], c["demand"], 1.0)
    ref = port.am_rollout("sdvrp", o0, h, W, "greedy")
    got = out["actions"].cpu()
    Tm = min(got.shape[1], ref["actions"].shape[1])
    same = (got[:, :Tm] == ref["actions"][:, :Tm]).all(1)
    assert same.float().mean() >= 0.9, "more than near-tie flips between the kernel's and the oracle's greedy tours"
    # log-probs of the ORACLE's tours through the evaluate path: no argmax divergence possible
    ev = pol(td, env, actions=ref["actions"].to(dev), hidden=h.to(dev), return_sum_log_likelihood=False)
    torch.testing.assert_close(ev["log_likelihood"].cpu(), ref["logprobs"], rtol=1e-4, atol=5e-5)
    torch.testing.assert_close(ev["reward"].cpu(), ref["reward"], rtol=RTOL, atol=2e-6)
    rows = same.nonzero()[:, 0]
    torch.testing.assert_close(out["reward"].cpu()[rows], ref["reward"][rows], rtol=RTOL, atol=2e-6)


def test_sampling_is_feasible_and_consistent(dev):
    torch.manual_seed(13)
    env = R.SDVRPEnv(generator_params=dict(num_loc=50), device=dev)
    pol = R.AttentionModelPolicy(env_name="sdvrp", num_encoder_layers=1).to(dev).eval()
    td = env.reset(env.generator(512))
    out = pol(td, env, decode_type="sampling", return_actions=True, return_hidden=True)
    env.check_solution_validity(td, out["actions"])
    c = {k: v.cpu() for k, v in td.items()}
    port.sdvrp_check(c["demand"], c["vehicle_capacity"], out["actions"].cpu())
    ev = pol(td, env, actions=out["actions"], hidden=out["hidden"].node_embeddings)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=RTOL, atol=5e-5)
    torch.testing.assert_close(ev["reward"], out["reward"], rtol=RTOL, atol=2e-6)
    torch.testing.assert_close(out["reward"].cpu(), port.cvrp_reward(c["locs"], out["actions"].cpu()), rtol=RTOL, atol=2e-6)


def _weights(pol):
    sd = {k: v.detach().cpu() for k, v in pol.state_dict().items()}
    return {"Wnode": sd["decoder.project_node_embeddings.weight"], "Wfixed": sd["decoder.project_fixed_context.weight"],
            "Wout": sd["decoder.pointer.project_out.weight"], "Wctx": sd["decoder.context_embedding.project_context.weight"],
            "Wdyn": sd["decoder.dynamic_embedding.projection.weight"]}


@pytest.mark.parametrize("size,B,S", [(20, 12, None), (50, 6, 17)])
def test_multistart_against_the_oracle(size, B, S, dev):
    """dynamic embedding x starts (am/decoder.py:176-187): keys and values are not shared across starts, every start is
    an independent episode that begins at customer s % num_loc + 1 with log-prob 0"""
    torch.manual_seed(500 + size)
    env = R.SDVRPEnv(generator_params=dict(num_loc=size), device=dev)
    pol = R.AttentionModelPolicy(env_name="sdvrp", num_encoder_layers=1).to(dev).eval()
    with torch.no_grad():
        pol.decoder.dynamic_embedding.projection.weight.mul_(3.0)
    td0 = env.generator(B)
    td = env.reset(td0)
    kw = {} if S is None else {"num_starts": S}
    out = pol(td, env, decode_type="multistart_greedy", return_actions=True, return_hidden=True,
              return_sum_log_likelihood=False, **kw)
    Sx = size if S is None else S
    assert out["reward"].shape == (B * Sx,)
    first = out["actions"][:, 0].view(Sx, B)
    assert (first == (torch.arange(Sx, device=dev) % size + 1)[:, None]).all()
    assert (out["log_likelihood"][:, 0] == 0).all()
    h = out["hidden"].node_embeddings.cpu()
    c = td0.cpu()
    ref = port.am_rollout("sdvrp", port.sdvrp_reset(c["depot"], c["locs"], c["demand"], 1.0), h, _weights(pol),
                          "multistart_greedy", num_starts=Sx)
    got = out["actions"].cpu()
    Tm = min(got.shape[1], ref["actions"].shape[1])
    same = (got[:, :Tm] == ref["actions"][:, :Tm]).all(1) & (got[:, Tm:] == 0).all(1) & (ref["actions"][:, Tm:] == 0).all(1)
    assert same.float().mean() >= 0.9, "more than near-tie flips between the kernel's and the oracle's greedy tours"
    rows = same.nonzero()[:, 0]
    torch.testing.assert_close(out["reward"].cpu()[rows], ref["reward"][rows], rtol=RTOL, atol=2e-6)
    torch.testing.assert_close(out["log_likelihood"].cpu()[rows].sum(1), ref["log_likelihood"][rows], rtol=1e-4, atol=5e-5)
    # every start is the single-start episode of its instance with a given first action: the evaluate path on the tiled
    # batch reproduces rewards and the log-probs of all steps after the forced one
    ev = pol(R.batchify(td, Sx), env, actions=out["actions"], hidden=R.batchify(out["hidden"].node_embeddings, Sx),
             return_sum_log_likelihood=False)
    torch.testing.assert_close(ev["reward"], out["reward"], rtol=RTOL, atol=2e-6)
    torch.testing.assert_close(ev["log_likelihood"][:, 1:], out["log_likelihood"][:, 1:], rtol=1e-4, atol=5e-5)
    # sampled starts: feasible split deliveries
    smp = pol(td, env, decode_type="multistart_sampling", return_actions=True, **kw)
    env.check_solution_validity(R.batchify(td, Sx), smp["actions"])
    port.sdvrp_check(port.batchify(td["demand"].cpu(), Sx), port.batchify(td["vehicle_capacity"].cpu(), Sx), smp["actions"].cpu())
