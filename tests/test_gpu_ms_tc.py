"""The tensor-core multistart kernel (csrc/am_rollout_ms_tc.cu: starts = TMEM lanes, FP16 two-term split) against the
register-file SIMT kernel and the oracle: every size class (N = 2 ... 112 nodes, S = 1 ... 128 starts, fewer / more
instances than SMs, with and without graph context), greedy / sampling / evaluate, tsp and cvrp.
Multistart semantics (utils/decoding.py:265-313): start s of instance b is rollout s*B + b, its first action is forced to
node s % num_loc (+ 1 past a depot) with log-prob 0.  The oracle scores a given tour single-start, so its step-0 log-prob
(the forced action under the policy) is skipped in the comparisons."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from rl4co_slap_b200 import _lib
from tests.helpers import (LOGP_ATOL, LOGP_RTOL, assert_divergences_are_near_ties, load_policy, oracle_weights)

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _setup(name, num_loc, B, dev, seed, variant=None):
    torch.manual_seed(seed)
    pol = load_policy(name, dev, variant)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=num_loc))
    gen = env.generator(B)
    td = env.reset(gen.to(dev))
    with torch.inference_mode():
        h = pol.encoder(td)[0].float().contiguous()
    return pol, env, gen, td, h


def _oracle_state(name, gen, S):
    g = {k: gen[k].cpu() for k in gen.keys()}
    st = port.tsp_reset(g["locs"]) if name == "tsp" else port.cvrp_reset(g["depot"], g["locs"], g["demand"], 1.0)
    return {k: (port.batchify(v, S) if isinstance(v, torch.Tensor) else v) for k, v in st.items()}


def _with_impl(impl, fn):
    _lib.call("rl4co_set_rollout_impl", impl)
    try:
        return fn()
    finally:
        _lib.call("rl4co_set_rollout_impl", 0)


@pytest.mark.parametrize("name,num_loc,B,S,variant", (
    ("tsp", 2, 3, 2, None), ("tsp", 5, 7, 5, None), ("tsp", 16, 150, 16, None), ("tsp", 33, 40, 33, None),
    ("tsp", 64, 20, 64, None), ("tsp", 100, 149, 100, None), ("tsp", 112, 9, 112, None), ("tsp", 50, 30, 128, None),
    ("cvrp", 2, 5, 2, None), ("cvrp", 10, 160, 10, None), ("cvrp", 31, 33, 31, None), ("cvrp", 63, 12, 63, None),
    ("cvrp", 100, 150, 100, "cvrp_pomo"), ("cvrp", 111, 6, 111, None), ("cvrp", 20, 64, 1, None), ("cvrp", 40, 10, 128, None)))
def test_ms_tc_greedy_and_sampling_against_simt_and_oracle(name, num_loc, B, S, variant, dev):
    pol, env, gen, td, h = _setup(name, num_loc, B, dev, 1000 + num_loc + S, variant)
    kw = dict(num_starts=S, hidden=h, return_actions=True, return_sum_log_likelihood=False)
    with torch.inference_mode():
        tc = pol(td, env, decode_type="multistart_greedy", **kw)
        simt = _with_impl(2, lambda: pol(td, env, decode_type="multistart_greedy", **kw))
        torch.manual_seed(5)
        smp = pol(td, env, decode_type="multistart_sampling", **kw)
    N = h.shape[1]
    tdS = R.batchify(td, S)
    first = torch.arange(S, device=dev).repeat_interleave(B) % num_loc + (name != "tsp")
    for out in (tc, smp):
        assert out["actions"].shape[0] == S * B and torch.equal(out["actions"][:, 0], first)
        assert (out["log_likelihood"][:, 0] == 0).all()
        env.check_solution_validity(tdS, out["actions"])
        torch.testing.assert_close(env.get_reward(tdS, out["actions"]), out["reward"], rtol=RTOL, atol=1e-5)
    # tensor-core kernel vs SIMT kernel: same tours except verified near ties
    W = oracle_weights(name, variant)
    ug = {} if variant != "cvrp_pomo" else {"use_graph_context": False}
    st, hS = _oracle_state(name, gen, S), R.batchify(h, S).cpu()
    # every diverging tour must split at a tie of the oracle's log-probs tighter than 1e-5 (observed: <= 1.5e-6); the
    # starts of an instance share partial tours, so one tie can flip several of them at once
    n_div, gap = assert_divergences_are_near_ties(name, st, hS, W, simt["actions"], tc["actions"], eps=1e-5, **ug)
    Tm = min(tc["actions"].shape[1], simt["actions"].shape[1])
    same = (tc["actions"][:, :Tm] == simt["actions"][:, :Tm]).all(1)
    assert n_div <= max(2, int(0.03 * S * B)), f"{n_div} of {S * B} tours flip at near ties (largest gap {gap:.2e})"
    torch.testing.assert_close(tc["log_likelihood"][:, :Tm][same], simt["log_likelihood"][:, :Tm][same], rtol=LOGP_RTOL, atol=LOGP_ATOL)
    torch.testing.assert_close(tc["reward"][same], simt["reward"][same], rtol=RTOL, atol=1e-5)
    # both tour sets against the oracle on a sub-sample (evaluate mode: no arg-max involved), steps 1 ...
    idx = torch.arange(0, S * B, max(1, S * B // 24))
    sub = {k: (v[idx] if isinstance(v, torch.Tensor) and v.dim() > 0 and v.shape[0] == S * B else v) for k, v in st.items()}
    for out in (tc, smp):
        acts = out["actions"][idx.to(dev)].cpu()
        ref = port.am_rollout(name, sub, hS[idx], W, actions=acts, check_solution=False, **ug)
        Tr = ref["logprobs"].shape[1]
        got = out["log_likelihood"][idx.to(dev)].cpu()[:, 1:Tr].double()
        want = ref["logprobs"][:, 1:].double()
        err, bound = (got - want).abs(), LOGP_RTOL * want.abs() + LOGP_ATOL
        assert (err <= bound).all(), f"per-step log-prob off by {err.max():.3e}"
        ll_got, ll_want = got.sum(1), want.sum(1)
        assert ((ll_got - ll_want).abs() <= RTOL * ll_want.abs() + 2e-5).all()
        torch.testing.assert_close(out["reward"][idx.to(dev)].cpu(), ref["reward"], rtol=RTOL, atol=1e-5)
    if N > 4 and S * B > 8:
        assert not torch.equal(smp["actions"], tc["actions"]), "sampling reproduced the greedy tours"


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_ms_tc_evaluate_mode_through_the_c_abi(name, dev):
    """mode = EVALUATE with multistart set (only reachable through the C ABI): the kernel's own greedy tours fed back
    give the same per-step log-probs, rewards and step counts; a tour with a masked action raises the status bit"""
    B, S, num_loc = 21, 20, 20
    pol, env, gen, td, h = _setup(name, num_loc, B, dev, 77)
    N = h.shape[1]
    T = N if name == "tsp" else 2 * (N - 1) + 1
    W, q0, wcap, _ = pol.decoder.folded_weights()
    nt = W.shape[0] // 128
    stream = _lib.stream_ptr(dev)
    tables = torch.empty(nt, B * N, 128, device=dev)
    _lib.call("rl4co_am_project_tc", B * N, nt, _lib.ptr(h), _lib.ptr(pol.decoder.weight_image()), _lib.ptr(tables), stream)
    gctx = torch.empty(B, 128, device=dev)
    _lib.call("rl4co_am_graph_context", B, N, _lib.ptr(h), _lib.ptr(pol.decoder.folded_weights()[3]), _lib.ptr(gctx), stream)

    def run(mode, actions_in=None):
        o = {"actions": torch.full((S * B, T), -1, dtype=torch.int64, device=dev), "logp": torch.zeros(S * B, T, device=dev),
             "ll": torch.zeros(S * B, device=dev), "reward": torch.zeros(S * B, device=dev),
             "steps": torch.zeros(S * B, dtype=torch.int32, device=dev), "status": _lib.new_status(dev)}
        a = _lib.AmRolloutArgs()
        a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = _lib.ENV_IDS[name], mode, B, S, 1, num_loc
        a.N, a.T, a.n_tables = N, T, nt
        a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
        a.tables, a.gctx, a.q0, a.wcap = _lib.ptr(tables), _lib.ptr(gctx), _lib.ptr(q0), _lib.ptr(wcap)
        a.locs = _lib.ptr(td["locs"].contiguous())
        if name == "cvrp":
            a.demand, a.vehicle_capacity = _lib.ptr(td["demand"].contiguous()), _lib.ptr(td["vehicle_capacity"].contiguous())
        a.actions_in = _lib.ptr(actions_in) if actions_in is not None else None
        a.actions, a.logp, a.ll, a.reward = _lib.ptr(o["actions"]), _lib.ptr(o["logp"]), _lib.ptr(o["ll"]), _lib.ptr(o["reward"])
        a.steps, a.status = _lib.ptr(o["steps"]), _lib.ptr(o["status"])
        _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
        torch.cuda.synchronize()
        return o

    g = run(_lib.MODE_GREEDY)
    assert int(g["status"].item()) == 0 and (g["actions"] >= 0).all()
    e = run(_lib.MODE_EVALUATE, g["actions"].contiguous())
    assert int(e["status"].item()) == 0
    assert torch.equal(e["actions"], g["actions"]) and torch.equal(e["steps"], g["steps"])
    torch.testing.assert_close(e["logp"], g["logp"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(e["reward"], g["reward"], rtol=0, atol=0)
    bad = g["actions"].clone()
    bad[3, 2] = bad[3, 1]                       # revisit a node: masked at that step
    e = run(_lib.MODE_EVALUATE, bad.contiguous())
    assert int(e["status"].item()) & (_lib.ST_INFEASIBLE_ACTION | _lib.ST_NEG_INF_LOGP)


def test_ms_tc_sampling_follows_the_policy_distribution(dev):
    """one TSP20 instance, start 3, 8192 copies: the first SAMPLED action's frequencies against the oracle's probabilities
    (chi-square), and the mean sampled reward within 4 standard errors of the oracle's own multinomial sampling"""
    from tests.helpers import fixture, t
    fx = fixture("tsp20")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    reps, S, s0 = 4096, 4, 3
    locs1, h1 = t(fx["gen.locs"])[:1], t(fx["hidden"])[:1]
    td = env.reset(R.TensorDict({"locs": locs1.expand(reps, -1, -1).contiguous().to(dev)}, batch_size=[reps]))
    hid = h1.expand(reps, -1, -1).contiguous().to(dev)
    torch.manual_seed(9)
    out = pol(td, env, decode_type="multistart_sampling", num_starts=S, return_actions=True, hidden=hid)
    acts = out["actions"].view(S, reps, -1)[s0].cpu()
    assert (acts[:, 0] == s0).all()
    W = oracle_weights("tsp")
    ref = port.am_rollout("tsp", port.tsp_reset(locs1), h1, W, actions=torch.tensor([[s0, 0]]), return_all_logp=True,
                          check_solution=False, max_steps=1)
    p1 = ref["logprobs"][0, 1].exp().double()
    counts = torch.bincount(acts[:, 1], minlength=20).double()
    keep = p1 * reps > 5
    chi2 = float((((counts - reps * p1) ** 2) / (reps * p1).clamp_min(1e-9))[keep].sum())
    assert chi2 < 40 + 3 * int(keep.sum()), chi2
    assert counts[s0] == 0
    torch.manual_seed(2)
    n_ref = 512
    o = port.tsp_reset(locs1.expand(n_ref, -1, -1).contiguous())
    refs = port.am_rollout("tsp", o, h1.expand(n_ref, -1, -1), W, "multistart_sampling", num_starts=S)
    a, b = out["reward"].double().cpu(), refs["reward"].double()
    se = (a.var() / a.numel() + b.var() / b.numel()).sqrt().item()
    assert abs(a.mean().item() - b.mean().item()) < 4 * se
