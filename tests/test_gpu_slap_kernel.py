"""The SLAP rollout kernel (csrc/am_rollout_slap.cu: constant query, glimpse maintained incrementally, node-major
ownership) against the oracle and against the generic register-file kernel (csrc/am_rollout_rf.cu), which recomputes the
whole attention at every step.  Sizes cover the four template instantiations (NW = 4, 8, 13, 16 nodes per warp); the
scaled encoder outputs make every head's attention collapse onto one or two locations, so that assigning them leaves a
sliver of the softmax denominator and the kernel's rescaling path (rebuild of all heads from the stored scores) runs -
without it the logits of the following steps would be cancellation noise and the comparison fails."""
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import assert_divergences_are_near_ties, load_policy, oracle_weights

RTOL = 1e-5                                       # BASELINE.json north_star tolerance on log-likelihoods and rewards

pytestmark = pytest.mark.gpu

SIZES = ((3, 7), (5, 10), (10, 10), (8, 16))      # N = 21, 50, 100, 128 storage locations


def _impl(k):
    from rl4co_slap_b200 import _lib
    _lib.call("rl4co_set_rollout_impl", k)


def _setup(dev, n_aisles, n_locs, B, scale, seed=5):
    torch.manual_seed(seed)
    env = R.SLAPEnv(generator_params={"n_aisles": n_aisles, "n_locs": n_locs}, device=dev)
    pol = load_policy("slap", dev)
    gen = env.generator(B)
    td = env.reset(gen.to(dev))
    with torch.inference_mode():
        h = pol.encoder(td)[0].float()
        # every generated instance has the same warehouse grid and therefore the same encoder output; the rollout takes any
        # `hidden`, so give every instance its own: per-node noise of a quarter of the embeddings' spread
        h = (h + 0.25 * h.std() * torch.randn_like(h)) * scale
    return env, pol, gen, td, h.contiguous()


def _oracle_state(gen, rows, P):
    cpu = {k: gen[k][rows].cpu() for k in ("locs", "freq", "assignment", "picklist", "depot_loc_dist")}
    return port.slap_reset(cpu["locs"], cpu["assignment"].clone(), P, {k: cpu[k] for k in ("freq", "picklist", "depot_loc_dist")})


def _oracle_ll(gen, h, acts, P, double):
    W = oracle_weights("slap")
    if double:
        W, h = {k: v.double() for k, v in W.items()}, h.double()
    return port.am_rollout("slap", _oracle_state(gen, slice(None), P), h, W, actions=acts,
                           check_solution=False)["log_likelihood"].double()


def _assert_fp32_class(got, ref32, ref64, other=None):
    """``got`` is held to RTOL |ll| + 2e-5 against the fp64 oracle, widened by what fp32 itself costs on this input: 4 x the
    fp32 oracle's own distance from fp64, and - where given - twice the distance of the generic kernel (both kernels read
    the same 3xTF32-projected tables, whose products carry 2^-21 rather than 2^-24).  The fixture policy on generated
    warehouses is badly conditioned: |h| up to 33 and more when scaled, glimpse scores in the hundreds."""
    err = (got - ref64).abs()
    bound = RTOL * ref64.abs() + 2e-5 + 4 * (ref32 - ref64).abs()
    if other is not None:
        bound = torch.maximum(bound, 2 * (other - ref64).abs() + 2e-5)
    assert (err <= bound).all(), (float(err.max()), float((err - bound).max()))


@pytest.mark.parametrize("n_aisles,n_locs", SIZES)
@pytest.mark.parametrize("scale", (1.0, 3.0, 6.0))
def test_against_the_generic_kernel(dev, n_aisles, n_locs, scale):
    B = 192
    env, pol, gen, td, h = _setup(dev, n_aisles, n_locs, B, scale)
    P = td["freq"].shape[-2]
    out = {}
    try:
        for impl in (0, 2):
            _impl(impl)
            with torch.inference_mode():
                out[impl] = {"greedy": pol(td, env, decode_type="greedy", hidden=h, return_actions=True)}
                torch.manual_seed(11)       # the Philox seed of a sampled rollout is drawn from torch's generator
                out[impl]["sampling"] = pol(td, env, decode_type="sampling", hidden=h, return_actions=True)
                out[impl]["evaluate"] = pol(td, env, actions=out[2 if 2 in out else 0]["sampling"]["actions"], hidden=h,
                                            return_actions=True)
    finally:
        _impl(0)
    W = oracle_weights("slap")
    a, b = out[0]["greedy"], out[2]["greedy"]
    assert a["actions"].shape == (B, P)
    # a different location is only ever picked where the oracle's two best candidates are an fp32 rounding apart
    n, _ = assert_divergences_are_near_ties("slap", _oracle_state(gen, slice(None), P), h, W, a["actions"], b["actions"])
    assert n <= max(2, B // 20)
    assert torch.equal(a["assignment"].long(), a["actions"])
    # same seed, same noise: sampled assignments agree except at such ties
    same = (out[0]["sampling"]["actions"] == out[2]["sampling"]["actions"]).all(1)
    assert int((~same).sum()) <= max(2, B // 20)
    torch.testing.assert_close(out[0]["sampling"]["reward"][same], out[2]["sampling"]["reward"][same], rtol=RTOL, atol=1e-5)
    # log-likelihoods against the fp64 oracle, next to the generic kernel's on the same assignments (greedy: the instances
    # where both kernels chose the same; evaluate: the generic kernel's sampled assignments given to both)
    assert torch.equal(out[0]["evaluate"]["actions"], out[2]["evaluate"]["actions"])
    rows_g = (a["actions"] == b["actions"]).all(1).cpu()
    for key, acts, rows in (("greedy", a["actions"].cpu(), rows_g), ("evaluate", out[0]["evaluate"]["actions"].cpu(), slice(None))):
        ref32, ref64 = _oracle_ll(gen, h.cpu(), acts, P, False), _oracle_ll(gen, h.cpu(), acts, P, True)
        _assert_fp32_class(out[0][key]["log_likelihood"].cpu().double()[rows], ref32[rows], ref64[rows],
                           out[2][key]["log_likelihood"].cpu().double()[rows])
    torch.testing.assert_close(out[0]["evaluate"]["reward"], out[2]["evaluate"]["reward"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("n_aisles,n_locs,scale", ((10, 10, 1.0), (10, 10, 3.0), (10, 10, 20.0), (5, 10, 4.0), (8, 16, 8.0),
                                                   (3, 7, 6.0)))
def test_given_assignments_against_the_oracle(dev, n_aisles, n_locs, scale):
    """evaluate mode: random feasible assignments - with collapsed attention (scale >= 3) many of them take a head's
    dominant location away and force the rebuild path - against oracle/port.py in fp32 and in fp64"""
    B = 64
    env, pol, gen, td, h = _setup(dev, n_aisles, n_locs, B, scale, seed=9)
    P, N = td["freq"].shape[-2], n_aisles * n_locs
    g = torch.Generator().manual_seed(3)
    acts = torch.stack([torch.randperm(N - 1, generator=g)[:P] + 1 for _ in range(B)])
    res = {}
    try:
        for impl in (0, 2):
            _impl(impl)
            with torch.inference_mode():
                res[impl] = pol(td, env, actions=acts.to(dev), hidden=h, return_actions=True)
    finally:
        _impl(0)
    ref32, ref64 = _oracle_ll(gen, h.cpu(), acts, P, False), _oracle_ll(gen, h.cpu(), acts, P, True)
    _assert_fp32_class(res[0]["log_likelihood"].cpu().double(), ref32, ref64, res[2]["log_likelihood"].cpu().double())
    ref = port.am_rollout("slap", _oracle_state(gen, slice(None), P), h.cpu(), oracle_weights("slap"), actions=acts)
    torch.testing.assert_close(res[0]["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-5)
    assert torch.equal(res[0]["actions"].cpu(), acts)


def test_repeated_location_is_reported(dev):
    env, pol, gen, td, h = _setup(dev, 10, 10, 8, 1.0)
    P = td["freq"].shape[-2]
    acts = torch.arange(1, P + 1).repeat(8, 1)
    acts[3, 7] = acts[3, 2]
    with pytest.raises(Exception, match="infeasible|Invalid|invalid|mask"):
        pol(td, env, actions=acts.to(dev), hidden=h)


def test_two_launches_agree_and_do_not_depend_on_the_batch_position(dev):
    """persistent CTAs carry registers / shared memory from one instance to the next (logit keys, value table, small
    inputs are fetched one instance ahead): an instance must come out the same wherever it sits in the batch"""
    env, pol, gen, td, h = _setup(dev, 10, 10, 1500, 1.0, seed=13)
    with torch.inference_mode():
        full = pol(td, env, decode_type="greedy", hidden=h, return_actions=True)
        again = pol(td, env, decode_type="greedy", hidden=h, return_actions=True)
        perm = torch.randperm(1500, device=dev)
        shuf = pol(td[perm], env, decode_type="greedy", hidden=h[perm], return_actions=True)
    assert torch.equal(full["actions"], again["actions"])
    assert torch.equal(full["log_likelihood"], again["log_likelihood"])
    assert torch.equal(full["actions"][perm], shuf["actions"])
    assert torch.equal(full["log_likelihood"][perm], shuf["log_likelihood"])
    assert torch.equal(full["reward"][perm], shuf["reward"])


@pytest.mark.parametrize("n_aisles,n_locs,P,O,K,graph_context", ((37, 1, 20, 20, 5, True), (8, 16, 32, 64, 4, True),
                                                                 (10, 10, 20, 20, 5, False), (2, 3, 5, 3, 2, True),
                                                                 (10, 10, 33, 20, 5, True)))
def test_limits_of_the_fixed_layout_and_odd_shapes(dev, n_aisles, n_locs, P, O, K, graph_context):
    """N = 37 (a prime), the largest shape the kernel's fixed shared-memory layout takes (N = 128, P = 32 steps, 256 picklist
    entries), no graph context (the query is zero: uniform attention), a tiny warehouse, and P = 33 - one step beyond the
    layout, which must fall through to the generic kernel with the same results: greedy assignments and evaluate
    log-likelihoods against the oracle"""
    B = 48
    torch.manual_seed(31)
    env = R.SLAPEnv(generator_params={"n_aisles": n_aisles, "n_locs": n_locs, "n_products": P, "max_orders": O,
                                      "max_products_in_order": K}, device=dev)
    pol = R.AttentionModelPolicy(env_name="slap", num_encoder_layers=1, feedforward_hidden=128,
                                 use_graph_context=graph_context).to(dev).eval()
    sd = {k: v for k, v in __import__("tests.helpers", fromlist=["policy_state"]).policy_state("slap").items()
          if not (k.endswith("project_fixed_context.weight") and not graph_context)}
    pol.load_state_dict(sd, strict=False)
    gen = env.generator(B)
    td = env.reset(gen.to(dev))
    with torch.inference_mode():
        h = pol.encoder(td)[0].float().contiguous()
        out = pol(td, env, decode_type="greedy", hidden=h, return_actions=True)
    W = oracle_weights("slap")
    state = _oracle_state(gen, slice(None), P)
    ref = port.am_rollout("slap", state, h.cpu(), W, "greedy", use_graph_context=graph_context)
    # (every instance of a generated batch has the same warehouse grid, hence the same encoder output: a verified near tie is the
    # same near tie in all of them, so the NUMBER of diverging instances says nothing here)
    assert_divergences_are_near_ties("slap", _oracle_state(gen, slice(None), P), h, W, out["actions"], ref["actions"],
                                     use_graph_context=graph_context)
    same = (out["actions"].cpu() == ref["actions"]).all(1)
    torch.testing.assert_close(out["reward"].cpu()[same], ref["reward"][same], rtol=RTOL, atol=1e-5)
    with torch.inference_mode():
        ev = pol(td, env, actions=ref["actions"].to(dev), hidden=h)
    ref64 = port.am_rollout("slap", _oracle_state(gen, slice(None), P), h.cpu().double(), {k: v.double() for k, v in W.items()},
                            actions=ref["actions"], use_graph_context=graph_context, check_solution=False)["log_likelihood"]
    _assert_fp32_class(ev["log_likelihood"].cpu().double(), ref["log_likelihood"].double(), ref64)
    torch.testing.assert_close(ev["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-5)
