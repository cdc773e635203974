"""tsp without table 4: the rollout kernels form the first node's context row themselves (args h / w_first, ABI 4).

The row is project_context.weight[:, :E] . h_first (nn/env_embeddings/context.py:123-131); K3 used to write it for every
node although a trajectory reads one.  Checked under all three kernel selections (the compaction kernel and the shared-memory
kernel form the row; with the compaction kernel switched off the call skips the register-file kernel, which does not take
h / w_first), in all three decode modes, against the five-table path and against the oracle in fp64.
Tolerances: the suite's 1e-5 relative on the log-likelihood (+2e-5 absolute, the fp32 noise floor of a 100-step sum);
tours of the two paths may split only at verified near ties of the oracle's log-probs.
"""

import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib
from oracle import port
from tests.helpers import assert_divergences_are_near_ties, load_policy, oracle_weights

pytestmark = pytest.mark.gpu
RTOL = 1e-5
IMPLS = {"auto": 0, "smem": 1, "rf": 2}


def _setup(B, N, dev, seed=0):
    torch.manual_seed(seed)
    pol = load_policy("tsp", dev)
    env = R.get_env("tsp", device=dev, generator_params=dict(num_loc=N))
    td = env.reset(batch_size=[B])
    with torch.inference_mode():
        h = pol.encoder(td)[0].float().contiguous()
    return pol, env, td, h


def _call(pol, td, env, h, in_kernel, impl="auto", **kw):
    pol.first_context_in_kernel = "always" if in_kernel else "never"
    _lib.call("rl4co_set_rollout_impl", IMPLS[impl])
    try:
        with torch.inference_mode():
            return pol(td, env, hidden=h, return_actions=True, **kw)
    finally:
        _lib.call("rl4co_set_rollout_impl", 0)
        pol.first_context_in_kernel = "auto"


@pytest.mark.parametrize("impl", list(IMPLS))
@pytest.mark.parametrize("N", [5, 20, 50, 100])
def test_evaluate_matches_five_tables_and_oracle(impl, N):
    dev = torch.device("cuda:0")
    B = 96
    pol, env, td, h = _setup(B, N, dev, seed=N)
    acts = torch.stack([torch.randperm(N, device=dev) for _ in range(B)])
    five = _call(pol, td, env, h, False, impl, actions=acts)
    four = _call(pol, td, env, h, True, impl, actions=acts)
    assert pol.last_launches >= 2
    torch.testing.assert_close(four["log_likelihood"], five["log_likelihood"], rtol=RTOL, atol=2e-5)
    assert torch.equal(four["reward"], five["reward"])
    W = {k: v.double() for k, v in oracle_weights("tsp").items()}
    ref = port.am_rollout("tsp", port.tsp_reset(td["locs"].cpu()), h.cpu().double(), W, actions=acts.cpu(),
                          check_solution=False)
    torch.testing.assert_close(four["log_likelihood"].cpu().double(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)


@pytest.mark.parametrize("impl", list(IMPLS))
def test_greedy_tours_split_only_at_near_ties(impl):
    dev = torch.device("cuda:0")
    B, N = 2048, 100
    pol, env, td, h = _setup(B, N, dev, seed=7)
    five = _call(pol, td, env, h, False, impl, decode_type="greedy")
    four = _call(pol, td, env, h, True, impl, decode_type="greedy")
    same = (five["actions"] == four["actions"]).all(1)
    assert same.float().mean().item() >= 0.99
    n_div, _ = assert_divergences_are_near_ties("tsp", port.tsp_reset(td["locs"].cpu()), h, oracle_weights("tsp"),
                                                five["actions"], four["actions"])
    assert n_div == int((~same).sum())
    assert torch.equal(four["actions"].sort(1).values, torch.arange(N, device=dev).expand(B, N))
    _, r = env.replay(td, four["actions"], check=True)
    torch.testing.assert_close(r, four["reward"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(four["log_likelihood"][same], five["log_likelihood"][same], rtol=RTOL, atol=2e-5)


@pytest.mark.parametrize("impl", ["auto", "smem"])
def test_sampling_same_noise_same_tours(impl):
    """Same Philox seed -> same noise: the two paths draw the same tours except at near ties.  (Not with the "rf" selection:
    there the five-table call samples on the register-file kernel and the four-table call on the shared-memory kernel, which
    draw differently.)"""
    dev = torch.device("cuda:0")
    B, N = 1024, 50
    pol, env, td, h = _setup(B, N, dev, seed=11)
    torch.manual_seed(5)
    five = _call(pol, td, env, h, False, impl, decode_type="sampling")
    torch.manual_seed(5)
    four = _call(pol, td, env, h, True, impl, decode_type="sampling")
    same = (five["actions"] == four["actions"]).all(1)
    assert same.float().mean().item() >= 0.98
    torch.testing.assert_close(four["log_likelihood"][same], five["log_likelihood"][same], rtol=RTOL, atol=2e-5)
    ev = _call(pol, td, env, h, True, impl, actions=four["actions"])
    torch.testing.assert_close(ev["log_likelihood"], four["log_likelihood"], rtol=RTOL, atol=2e-5)


def test_all_log_probs_and_multistart_fall_back():
    """store_all_logp runs on the shared-memory kernel with four tables; multistart keeps five (the tensor-core kernel
    reads table 4 for every start) and return_hidden hands back all five tables."""
    dev = torch.device("cuda:0")
    B, N = 64, 20
    pol, env, td, h = _setup(B, N, dev, seed=3)
    acts = torch.stack([torch.randperm(N, device=dev) for _ in range(B)])
    pol.first_context_in_kernel = "never"
    with torch.inference_mode():
        five = pol(td, env, hidden=h, actions=acts, store_all_logp=True, return_entropy=True)
        pol.first_context_in_kernel = "always"
        four = pol(td, env, hidden=h, actions=acts, store_all_logp=True, return_entropy=True)
        torch.testing.assert_close(four["log_likelihood"], five["log_likelihood"], rtol=RTOL, atol=2e-5)
        torch.testing.assert_close(four["entropy"], five["entropy"], rtol=1e-4, atol=1e-5)
        ms = pol(td, env, hidden=h, decode_type="multistart_greedy", num_starts=N, return_actions=True)
        assert ms["reward"].shape[0] == B * N
        hid = pol(td, env, decode_type="greedy", return_hidden=True)
    assert hid["hidden"].tables.shape[0] == 5


def test_c_abi_argument_checks():
    dev = torch.device("cuda:0")
    pol, env, td, h = _setup(4, 10, dev)
    W, q0, _, Wfixed = pol.decoder.folded_weights()
    tables = torch.empty(4, 40, 128, device=dev)
    a = _lib.AmRolloutArgs()
    a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = 0, _lib.MODE_GREEDY, 4, 0, 0, 10
    a.N, a.T, a.n_tables = 10, 10, 4
    a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
    acts = torch.empty(4, 10, dtype=torch.int64, device=dev)
    ll, rew = torch.empty(4, device=dev), torch.empty(4, device=dev)
    a.tables, a.q0, a.locs = _lib.ptr(tables), _lib.ptr(q0), _lib.ptr(td["locs"].contiguous())
    a.actions, a.ll, a.reward, a.status = _lib.ptr(acts), _lib.ptr(ll), _lib.ptr(rew), _lib.ptr(_lib.new_status(dev))
    stream = _lib.stream_ptr(dev)
    with pytest.raises(RuntimeError, match="needs 5 tables"):
        _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
    a.h = _lib.ptr(h)
    with pytest.raises(RuntimeError, match="come together"):
        _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
    a.w_first = _lib.ptr(W[512:640])
    _lib.call("rl4co_am_project_tc", 40, 4, _lib.ptr(h), _lib.ptr(pol.decoder.weight_image()), _lib.ptr(tables), stream)
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
    torch.cuda.synchronize()
    assert torch.equal(acts.sort(1).values, torch.arange(10, device=dev).expand(4, 10))
    a.env = 1
    with pytest.raises(RuntimeError, match="tsp arguments"):
        _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
