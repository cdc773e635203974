"""GPU parity: env kernels (K1) through the reference-shaped API vs the reference fixtures
and the oracle port.  Masks / states / tours bit-exact; rewards within 1e-5 relative."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import (ATSP_CASES, CVRP_CASES, OP_CASES, SDVRP_CASES, PCTSP_CASES, SLAP_CASES, TSP_CASES, env_of, fixture, gen_td, oracle_reset, t)

pytestmark = pytest.mark.gpu


def _keys(fx, prefix):
    return [k[len(prefix) + 1:] for k in fx.files if k.startswith(prefix + ".")]


def _close(got, want, what):
    """rewards: 1e-5 relative (north_star); the kernels sum the legs warp-parallel, torch
    sums them in its own vectorised order, so the last ulp may differ"""
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-5, atol=1e-6, err_msg=str(what))
    assert got.cpu().numpy().dtype == want.dtype and got.shape == want.shape, what


# the one state tensor whose value depends on torch's own reduction order: pctsp's cur_total_penalty starts from
# penalty.sum(-1) (pctsp/env.py:123), which the reference itself computes differently on CPU and on CUDA
_SUM_ORDER_KEYS = ("cur_total_penalty",)


def _same(got, want, what):
    if what[-1] in _SUM_ORDER_KEYS:
        return _close(got, want, what)
    got = got.cpu().numpy()
    assert got.dtype == want.dtype, (what, got.dtype, want.dtype)
    assert got.shape == want.shape, (what, got.shape, want.shape)
    assert np.array_equal(got, want), what


@pytest.mark.parametrize("case", TSP_CASES + CVRP_CASES + SLAP_CASES + OP_CASES + PCTSP_CASES + ATSP_CASES + SDVRP_CASES)
def test_step_by_step_replay_bit_exact(case, dev):
    """reset + every _step through the env API: each state tensor equals the reference's
    (dtype, shape and value) after every step; then get_reward."""
    fx, name = fixture(case), env_of(case)
    env = R.get_env(name, device=dev)
    td = env.reset(gen_td(name, fx, dev))
    for k in _keys(fx, "reset"):
        _same(td[k], fx[f"reset.{k}"], (case, "reset", k))
    acts = t(fx["replay.actions"], dev)
    for step in range(int(fx["replay.steps"])):
        td.set("action", acts[:, step])
        td = env.step(td)["next"]
        for k in _keys(fx, f"step{step}"):
            _same(td[k], fx[f"step{step}.{k}"], (case, step, k))
    assert bool(td["done"].all())
    reward = env.get_reward(td, acts)
    _close(reward, fx["replay.reward"], (case, "reward"))


@pytest.mark.parametrize("case", TSP_CASES + CVRP_CASES + SLAP_CASES + OP_CASES + PCTSP_CASES + ATSP_CASES + SDVRP_CASES)
def test_fused_replay_matches_fixture(case, dev):
    fx, name = fixture(case), env_of(case)
    env = R.get_env(name, device=dev)
    td = env.reset(gen_td(name, fx, dev))
    acts = t(fx["replay.actions"], dev)
    final, reward = env.replay(td, acts)
    _close(reward, fx["replay.reward"], (case, "reward"))
    last = int(fx["replay.steps"]) - 1
    for k in _keys(fx, f"step{last}"):
        if k in final and k != "reward":
            _same(final[k], fx[f"step{last}.{k}"], (case, "final", k))


def _random_feasible(name, td0, gen):
    """random feasible action sequences from the oracle env (CPU)"""
    td = dict(td0)
    acts = []
    while not td["done"].all():
        a = torch.multinomial(td["action_mask"].float(), 1, generator=gen).squeeze(-1)
        td["action"] = a
        td = port._STEP[name](td)
        acts.append(a)
    return torch.stack(acts, 1), td


@pytest.mark.parametrize("name,size,B", [("tsp", 100, 257), ("tsp", 37, 100), ("cvrp", 100, 129),
                                         ("cvrp", 21, 64), ("slap", 100, 65)])
def test_fused_replay_vs_oracle_random(name, size, B, dev):
    gen = torch.Generator().manual_seed(1234 + size)
    if name == "tsp":
        g = R.TSPGenerator(num_loc=size)
    elif name == "cvrp":
        g = R.CVRPGenerator(num_loc=size)
    else:
        g = R.SLAPGenerator(n_aisles=size // 10, n_locs=10)
    torch.manual_seed(99 + size)
    td0 = g(B)
    if name == "tsp":
        o0 = port.tsp_reset(td0["locs"])
    elif name == "cvrp":
        o0 = port.cvrp_reset(td0["depot"], td0["locs"], td0["demand"], 1.0)
    else:
        o0 = port.slap_reset(td0["locs"], td0["assignment"], 20, {"picklist": td0["picklist"], "freq": td0["freq"]})
    acts, fin = _random_feasible(name, o0, gen)
    if name == "tsp":
        want = port.tsp_reward(fin["locs"], acts)
    elif name == "cvrp":
        port.cvrp_check(fin["demand"], fin["vehicle_capacity"], acts)
        want = port.cvrp_reward(fin["locs"], acts)
    else:
        want = port.slap_reward(fin["locs"], fin["assignment"], fin["picklist"])
    env = R.get_env(name, device=dev)
    td = env.reset(td0.to(dev))
    final, reward = env.replay(td, acts.to(dev))
    torch.testing.assert_close(reward.cpu(), want, rtol=1e-5, atol=1e-6)
    for k in ("action_mask", "visited", "used_capacity", "current_node", "assignment", "done"):
        if k in final and k in fin:
            assert torch.equal(final[k].cpu().view(fin[k].shape), fin[k]), k
    torch.testing.assert_close(env.get_reward(final if name == "slap" else td, acts.to(dev)).cpu(), want, rtol=1e-5, atol=1e-6)


def test_invalid_tours_raise_reference_messages(dev):
    fx = fixture("tsp20")
    env = R.TSPEnv(device=dev)
    td = env.reset(gen_td("tsp", fx, dev))
    bad = t(fx["replay.actions"], dev).clone()
    bad[3, 5] = bad[3, 4]
    with pytest.raises(AssertionError, match="Invalid tour"):
        env.get_reward(td, bad)
    env.check_solution = False
    env.get_reward(td, bad)  # no check -> no raise, like the reference
    fx = fixture("cvrp20")
    cenv = R.CVRPEnv(device=dev)
    td = cenv.reset(gen_td("cvrp", fx, dev))
    acts = t(fx["replay.actions"], dev)
    bad = acts.clone()
    bad[0, 0] = bad[0, 1] if bad[0, 1] != 0 else bad[0, 2]
    with pytest.raises(AssertionError, match="Invalid tour"):
        cenv.get_reward(td, bad)
    # all customers in one trip without returning: capacity violated but a valid permutation
    n = acts.shape[0]
    over = torch.arange(1, 21, device=dev).repeat(n, 1)
    with pytest.raises(AssertionError, match="Used more than capacity"):
        cenv.get_reward(td, over)
    with pytest.raises(AssertionError, match="Used more than capacity"):
        port.cvrp_check(t(fx["gen.demand"]), torch.ones(n, 1), over.cpu())


def test_empty_batch_and_bad_arguments(dev):
    env = R.TSPEnv(generator_params=dict(num_loc=10), device=dev)
    td = env.reset(batch_size=[0])
    assert td["action_mask"].shape == (0, 10)
    _, r = env.replay(td, torch.empty(0, 10, dtype=torch.int64, device=dev))
    assert r.shape == (0,)
    td = env.reset(batch_size=[3])
    with pytest.raises(RuntimeError, match="expected dtype"):
        env.replay(td, torch.zeros(3, 10, dtype=torch.int32, device=dev))
    with pytest.raises(AssertionError, match="Invalid tour"):
        env.replay(td, torch.full((3, 10), 10, dtype=torch.int64, device=dev))
    with pytest.raises(AssertionError, match="index out of range"):
        env.replay(td, torch.full((3, 10), 10, dtype=torch.int64, device=dev), check=False)


def test_tsp_first_step_is_batch_global(dev):
    """tsp/env.py:70: `td["i"].all() == 0` is a whole-batch test - mixed i values keep the
    reference's (quirky) behaviour."""
    env = R.TSPEnv(generator_params=dict(num_loc=6), device=dev)
    td = env.reset(batch_size=[4])
    td["i"][1:] = 3                      # i.all() is False -> `False == 0` -> first_node overwritten
    td.set("action", torch.tensor([2, 3, 4, 5], device=dev))
    o = port.tsp_reset(td["locs"].cpu())
    o["i"] = td["i"].cpu().clone()
    o["action"] = td["action"].cpu()
    o = port.tsp_step(o)
    td = env.step(td)["next"]
    assert torch.equal(td["first_node"].cpu(), o["first_node"])
    td["i"][:] = 2                       # all non-zero -> first_node kept
    keep = td["first_node"].clone()
    td.set("action", torch.tensor([0, 1, 0, 1], device=dev))
    td = env.step(td)["next"]
    assert torch.equal(td["first_node"], keep)


def test_env_rollout_with_random_policy_shapes(dev):
    """the reference's env smoke test (tests/test_envs.py:39-59): reward shape (B,)"""
    for name, kw in (("tsp", dict(num_loc=20)), ("cvrp", dict(num_loc=20)), ("slap", dict(n_aisles=3, n_locs=10))):
        env = R.get_env(name, generator_params=kw, device=dev)
        reward, td, actions = R.rollout(env, env.reset(batch_size=[2]), R.random_policy)
        assert reward.shape == (2,)


def test_full_size_replay_properties(dev):
    """BASELINE config 2 size (B = 2^20, TSP100): size-independent checks - rotating or
    reversing a tour keeps its length (to fp32 sum reordering), final mask is all zero, a
    sub-sample equals the oracle bit for bit."""
    B, N = 1 << 20, 100
    torch.manual_seed(1234)
    locs = torch.rand(B, N, 2, device=dev)
    acts = torch.rand(B, N, device=dev).argsort(1)
    env = R.TSPEnv(device=dev)
    td = R.TensorDict({"locs": locs}, batch_size=[B])
    final, r = env.replay(td, acts)
    assert not final["action_mask"].any() and final["done"].all()
    assert torch.equal(final["first_node"], acts[:, 0]) and torch.equal(final["current_node"], acts[:, -1])
    _, r_rot = env.replay(td, acts.roll(17, 1))
    _, r_rev = env.replay(td, acts.flip(1))
    torch.testing.assert_close(r_rot, r, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(r_rev, r, rtol=1e-5, atol=1e-5)
    idx = torch.randint(0, B, (512,), device=dev)
    want = port.tsp_reward(locs[idx].cpu(), acts[idx].cpu())
    torch.testing.assert_close(r[idx].cpu(), want, rtol=1e-5, atol=1e-6)
    assert abs(r.double().mean().item() + 52.14) < 0.05   # theory: N * 0.5214


@pytest.mark.parametrize("size,B", [(100, 300), (37, 64), (128, 33)])
def test_atsp_generator_and_replay_vs_oracle(size, B, dev):
    """cost matrices from the on-device generator satisfy the triangle inequality (the generator's fixed point);
    random permutations replayed through the fused kernel against the oracle's gather"""
    torch.manual_seed(31 + size)
    env = R.ATSPEnv(generator_params=dict(num_loc=size), device=dev)
    td0 = env.generator(B)
    cm = td0["cost_matrix"]
    assert cm.shape == (B, size, size) and bool((cm.diagonal(dim1=-2, dim2=-1) == 0).all())
    sub = cm[:8]
    closure = (sub[:, :, None, :] + sub[:, None, :, :].transpose(-2, -1)).min(-1)[0]
    assert torch.equal(closure, sub), "triangle inequality (min-plus fixed point) violated"
    td = env.reset(td0)
    acts = torch.rand(B, size, device=dev).argsort(1)
    final, reward = env.replay(td, acts)
    want = port.atsp_reward(cm.cpu(), acts.cpu())
    torch.testing.assert_close(reward.cpu(), want, rtol=1e-5, atol=1e-6)
    assert bool(final["done"].all()) and not bool(final["action_mask"].any())
    assert torch.equal(final["first_node"], acts[:, 0]) and torch.equal(final["current_node"], acts[:, -1])
    torch.testing.assert_close(env.get_reward(td, acts).cpu(), want, rtol=1e-5, atol=1e-6)
    bad = acts.clone()
    bad[1, 3] = bad[1, 2]
    with pytest.raises(AssertionError, match="Invalid tour"):
        env.check_solution_validity(td, bad)


def test_sdvrp_invalid_tours_raise_reference_messages(dev):
    fx = fixture("sdvrp20")
    env = R.SDVRPEnv(device=dev)
    td = env.reset(gen_td("sdvrp", fx, dev))
    acts = t(fx["replay.actions"], dev)
    env.check_solution_validity(td, acts)
    with pytest.raises(AssertionError, match="All demand must be satisfied"):
        env.check_solution_validity(td, acts[:, :5])
    bad = torch.cat((torch.zeros_like(acts[:, :2]), acts), 1)
    with pytest.raises(AssertionError, match="Cannot visit depot twice"):
        env.check_solution_validity(td, bad)
