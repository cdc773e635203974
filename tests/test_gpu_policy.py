"""GPU parity: projection (K3) + persistent rollout kernel (K2) + reductions (K4) through
AttentionModelPolicy.forward vs the reference fixtures and the oracle port.
Tolerances (BASELINE.json north_star): tours bit-exact given identical logits; rewards and
log-likelihoods within 1e-5 relative (fp32); sampling matches in distribution."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from oracle import port
from tests.helpers import (assert_divergences_are_near_ties, assert_logp_close, env_of, fixture, gen_td, load_policy,
                           oracle_reset, oracle_weights, t)

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(params=("auto", "rf", "smem"))
def rollout_impl(request):
    """Single-start rollouts have three kernels behind rl4co_am_rollout (register-file resident
    tables with node compaction for TSP greedy / register-file resident tables / shared-memory
    resident tables); the parity tests run all of them."""
    from rl4co_slap_b200 import _lib
    _lib.call("rl4co_set_rollout_impl", {"auto": 0, "smem": 1, "rf": 2}[request.param])
    yield request.param
    _lib.call("rl4co_set_rollout_impl", 0)


def _oracle_fp64(name, fx, actions):
    """The oracle's decoder evaluated in float64 on the same inputs (reference op order); the
    environment state stays fp32 so that the capacity masks are the reference's bit for bit."""
    td = oracle_reset(name, fx)
    W = {k: v.double() for k, v in oracle_weights(name).items()}
    return port.am_rollout(name, td, t(fx["hidden"]).double(), W, actions=actions, check_solution=False)


def _run(case, dev, variant=None, **kw):
    fx, name = fixture(case), env_of(case)
    pol = load_policy(name, dev, variant)
    env = R.get_env(name, device=dev)
    td = env.reset(gen_td(name, fx, dev))
    return fx, pol, env, td, pol(td, env, return_actions=True, hidden=t(fx["hidden"], dev), **kw)


@pytest.mark.parametrize("case", ("tsp5", "tsp20", "tsp50", "tsp100", "cvrp10", "cvrp20", "cvrp50", "cvrp100",
                                  "slap50", "slap100"))
def test_greedy_matches_reference_fixture(case, dev, rollout_impl):
    fx, pol, env, td, out = _run(case, dev, decode_type="greedy", return_sum_log_likelihood=False)
    assert out["actions"].dtype == torch.int64
    assert np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"]), "tours differ from the reference"
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["greedy.reward"], rtol=RTOL, atol=1e-6)
    # per step: helpers.LOGP_RTOL / LOGP_ATOL plus the fp32 reference's own distance from an fp64 run of itself
    noise = np.abs(_oracle_fp64(env_of(case), fx, t(fx["greedy.actions"]))["logprobs"].numpy() - fx["greedy.logprobs"])
    assert_logp_close(out["log_likelihood"].cpu().numpy(), fx["greedy.logprobs"], noise)
    np.testing.assert_allclose(out["log_likelihood"].sum(1).cpu().numpy(), fx["greedy.ll"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("case", ("tsp20", "tsp50", "cvrp20", "cvrp50"))
def test_multistart_greedy_matches_reference_fixture(case, dev):
    S = int(fixture(case)["ms_greedy.num_starts"])
    fx, pol, env, td, out = _run(case, dev, decode_type="multistart_greedy", num_starts=S)
    assert out["reward"].shape == (fx["gen.locs"].shape[0] * S,)      # tests/test_policy.py:37-51
    assert np.array_equal(out["actions"].cpu().numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["ms_greedy.reward"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["ms_greedy.ll"], rtol=RTOL, atol=1e-5)


def test_pomo_policy_multistart_default_starts(dev):
    fx, pol, env, td, out = _run("cvrp20_pomo", dev, variant="cvrp_pomo", decode_type="multistart_greedy")
    assert np.array_equal(out["actions"].cpu().numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["ms_greedy.ll"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("case", ("tsp50", "cvrp50", "slap100"))
def test_evaluate_given_actions(case, dev, rollout_impl):
    """ll of the reference's tours (no argmax divergence possible), constructive/base.py:206-207"""
    fx, name = fixture(case), env_of(case)
    pol, env = load_policy(name, dev), R.get_env(name, device=dev)
    td = env.reset(gen_td(name, fx, dev))
    out = pol(td, env, actions=t(fx["greedy.actions"], dev), hidden=t(fx["hidden"], dev), return_actions=True)
    assert np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"])
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), fx["greedy.ll"], rtol=RTOL, atol=1e-5)
    np.testing.assert_allclose(out["reward"].cpu().numpy(), fx["greedy.reward"], rtol=RTOL, atol=1e-6)
    # a random feasible tour against the oracle
    acts = t(fx["replay.actions"], dev)
    out = pol(td, env, actions=acts, hidden=t(fx["hidden"], dev))
    ref = port.am_rollout(name, oracle_reset(name, fx), t(fx["hidden"]), oracle_weights(name), actions=acts.cpu())
    # 1e-5 relative, plus the oracle's OWN fp32 rounding error on this input (its distance from an
    # fp64 run of the same oracle): a random tour through the slap fixture (|h| up to 33, logits up
    # to 31) has |ll| ~ 280 and the fp32 oracle itself is 4.5e-6 relative away from fp64 there
    ref64 = _oracle_fp64(name, fx, acts.cpu())
    noise = (ref["log_likelihood"].double() - ref64["log_likelihood"]).abs()
    err = (out["log_likelihood"].cpu().double() - ref["log_likelihood"].double()).abs()
    bound = RTOL * ref["log_likelihood"].double().abs() + 1e-5 + 4 * noise
    assert (err <= bound).all(), (err, bound)
    assert ((out["log_likelihood"].cpu().double() - ref64["log_likelihood"]).abs()
            <= RTOL * ref64["log_likelihood"].abs() + 1e-5 + 4 * noise).all()
    torch.testing.assert_close(out["reward"].cpu(), ref["reward"], rtol=RTOL, atol=1e-6)


def test_evaluate_invalid_tour_raises(dev):
    fx = fixture("tsp20")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    td = env.reset(gen_td("tsp", fx, dev))
    bad = t(fx["greedy.actions"], dev).clone()
    bad[0, 3] = bad[0, 2]
    with pytest.raises(AssertionError, match="Invalid tour"):
        pol(td, env, actions=bad, hidden=t(fx["hidden"], dev))


def test_full_policy_with_pytorch_encoder(dev):
    """encoder (PyTorch on GPU) + kernels end to end; larger random batch vs the oracle fed
    with the same GPU-computed embeddings"""
    torch.manual_seed(5)
    for name, g in (("tsp", R.TSPGenerator(num_loc=30)), ("cvrp", R.CVRPGenerator(num_loc=30))):
        pol, env = load_policy(name, dev), R.get_env(name, device=dev)
        td0 = g(96)
        td = env.reset(td0.to(dev))
        with torch.inference_mode():
            out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
        h = out["hidden"].node_embeddings.cpu()
        o0 = port.tsp_reset(td0["locs"]) if name == "tsp" else port.cvrp_reset(td0["depot"], td0["locs"], td0["demand"])
        ref = port.am_rollout(name, o0, h, oracle_weights(name), "greedy")
        ga, ra = out["actions"].cpu(), ref["actions"]
        Tm = max(ga.shape[1], ra.shape[1])      # cvrp: a near-tie flip may change the padded length
        ga = torch.nn.functional.pad(ga, (0, Tm - ga.shape[1]))
        ra = torch.nn.functional.pad(ra, (0, Tm - ra.shape[1]))
        same = (ga == ra).all(1)
        # every divergence must be a verified near tie at its first differing step (not merely "few of them")
        n_div, gap = assert_divergences_are_near_ties(name, o0, h, oracle_weights(name), ra, ga)
        assert n_div == int((~same).sum())
        assert same.float().mean() >= 0.97, f"{n_div} near-tie flips in 96 instances (largest gap {gap:.2e})"
        torch.testing.assert_close(out["reward"].cpu()[same], ref["reward"][same], rtol=RTOL, atol=1e-6)
        torch.testing.assert_close(out["log_likelihood"].cpu()[same], ref["log_likelihood"][same], rtol=RTOL, atol=2e-5)
        ev = pol(td, env, actions=ref["actions"].to(dev), hidden=h.to(dev))
        torch.testing.assert_close(ev["log_likelihood"].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)


def test_chunked_equals_unchunked(dev):
    fx = fixture("tsp50")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    td = env.reset(gen_td("tsp", fx, dev))
    a = pol(td, env, decode_type="greedy", return_actions=True)
    pol.rollout_chunk = 4
    b = pol(td, env, decode_type="greedy", return_actions=True)
    assert torch.equal(a["actions"], b["actions"]) and torch.equal(a["reward"], b["reward"])
    assert torch.equal(a["log_likelihood"], b["log_likelihood"])


def test_sampling_distribution(dev, rollout_impl):
    """Philox sampling vs the reference distribution: (1) every sampled tour is feasible and
    its ll equals the evaluate-mode ll of the same tour; (2) first-step action frequencies
    match the oracle's step-0 probabilities (chi-square); (3) mean reward of samples equals
    the oracle's own multinomial sampling within 4 standard errors."""
    fx = fixture("tsp20")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    h1 = t(fx["hidden"], dev)[:1]
    reps = 4096
    locs = t(fx["gen.locs"], dev)[:1].expand(reps, -1, -1).contiguous()
    td = env.reset(R.TensorDict({"locs": locs}, batch_size=[reps]))
    hid = h1.expand(reps, -1, -1).contiguous()
    torch.manual_seed(11)
    out = pol(td, env, decode_type="sampling", return_actions=True, hidden=hid)
    env.check_solution_validity(td, out["actions"])
    ev = pol(td, env, actions=out["actions"], hidden=hid)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=1e-6, atol=1e-6)
    again = pol(td, env, decode_type="sampling", return_actions=True, hidden=hid)
    assert not torch.equal(again["actions"], out["actions"])          # fresh seed per call
    # step-0 distribution from the oracle
    W = oracle_weights("tsp")
    o0 = port.tsp_reset(t(fx["gen.locs"])[:1])
    cache = port.precompute_cache(t(fx["hidden"])[:1], W)
    logits, mask = port.decoder_forward("tsp", o0, cache, W, 0)
    p0 = port.process_logits(logits, mask).exp()[0].double()
    counts = torch.bincount(out["actions"][:, 0].cpu(), minlength=20).double()
    chi2 = (((counts - reps * p0) ** 2) / (reps * p0).clamp_min(1e-9))[p0 * reps > 5].sum().item()
    assert chi2 < 60.0, chi2                                           # dof <= 19, p ~ 1e-6
    torch.manual_seed(3)
    o = port.tsp_reset(t(fx["gen.locs"])[:1].expand(1024, -1, -1).contiguous())
    ref = port.am_rollout("tsp", o, t(fx["hidden"])[:1].expand(1024, -1, -1), W, "sampling")
    a, b = out["reward"].double().cpu(), ref["reward"].double()
    se = (a.var() / a.numel() + b.var() / b.numel()).sqrt().item()
    assert abs(a.mean().item() - b.mean().item()) < 4 * se


def test_multistart_sampling_cvrp_pomo(dev):
    """BASELINE config 4 shape: CVRP POMO multistart sampling + shared baseline"""
    fx = fixture("cvrp20_pomo")
    pol, env = load_policy("cvrp", dev, "cvrp_pomo"), R.CVRPEnv(device=dev)
    td = env.reset(gen_td("cvrp", fx, dev))
    B, S = fx["gen.locs"].shape[0], 20
    torch.manual_seed(2)
    out = pol(td, env, decode_type="multistart_sampling", return_actions=True, hidden=t(fx["hidden"], dev))
    assert out["reward"].shape == (B * S,)
    first = out["actions"][:, 0].view(S, B)
    assert torch.equal(first, torch.arange(1, S + 1, device=dev)[:, None].expand(S, B))
    env.check_solution_validity(R.batchify(td, S), out["actions"])
    ref_r = port.cvrp_reward(R.batchify(td["locs"], S).cpu(), out["actions"].cpu())
    torch.testing.assert_close(out["reward"].cpu(), ref_r, rtol=RTOL, atol=1e-6)
    bl, part = R.SharedBaseline().eval_fused(out["reward"], out["log_likelihood"], num_starts=S)
    r2 = R.unbatchify(out["reward"], S)
    torch.testing.assert_close(bl, r2.mean(1, keepdim=True), rtol=1e-6, atol=1e-6)
    # the reference's surface (baselines.py:57-61): eval(td, reward[B,S], env) -> (mean over starts, 0); a flat start-major
    # reward works when the baseline knows S; through the registry / WarmupBaseline as reinforce_step calls it
    v, zero = R.SharedBaseline().eval(td, r2, env)
    assert zero == 0 and torch.equal(v, r2.mean(1, keepdim=True))
    v, _ = R.get_reinforce_baseline("shared", num_starts=S).eval(td, out["reward"], env)
    torch.testing.assert_close(R.unbatchify(v, S), r2.mean(1, keepdim=True).expand(-1, S), rtol=1e-6, atol=1e-6)
    wb = R.get_reinforce_baseline("warmup", baseline="shared", num_starts=S, n_epochs=2)
    wb.epoch_callback(epoch=0)      # alpha = 0.5: convex combination of the shared and the exponential baseline
    v, _ = wb.eval(td, out["reward"], env)
    want_v = 0.5 * r2.mean(1, keepdim=True).expand(-1, S) + 0.5 * out["reward"].mean()
    torch.testing.assert_close(R.unbatchify(v, S), want_v, rtol=1e-5, atol=1e-5)
    loss = R.reinforce_loss(part)
    # float64 reference; the fp32 baseline is one rounding (<= ulp(|r|)/2) away from the exact
    # mean and that offset multiplies every ll of the instance, hence the absolute tolerance
    ll2 = R.unbatchify(out["log_likelihood"], S).double()
    want = -((r2.double() - r2.double().mean(1, keepdim=True)) * ll2).mean()
    atol = 2 * float(torch.finfo(torch.float32).eps) * float(r2.abs().max()) * float(ll2.abs().mean())
    torch.testing.assert_close(loss, want.to(loss.dtype), rtol=1e-5, atol=atol)


def test_select_best_and_entropy(dev):
    fx = fixture("tsp20")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    td = env.reset(gen_td("tsp", fx, dev))
    h = t(fx["hidden"], dev)
    full = pol(td, env, decode_type="multistart_greedy", num_starts=8, return_actions=True, hidden=h)
    best = pol(td, env, decode_type="multistart_greedy", num_starts=8, return_actions=True, hidden=h,
               select_best=True, return_entropy=True)
    torch.testing.assert_close(best["reward"], R.unbatchify(full["reward"], 8).max(1)[0])
    assert best["actions"].shape == (8, 20) and best["entropy"].shape == (8,)
    ent = pol(td, env, decode_type="greedy", return_entropy=True, hidden=h)["entropy"]
    ref = port.am_rollout("tsp", oracle_reset("tsp", fx), t(fx["hidden"]), oracle_weights("tsp"), "greedy",
                          return_all_logp=True)
    lp = torch.nan_to_num(ref["logprobs"], nan=0.0)
    want = -(lp.exp() * lp).nan_to_num(0.0).sum(-1).sum(1)
    torch.testing.assert_close(ent.cpu(), want, rtol=1e-4, atol=1e-4)


def test_decode_step_vs_torch(dev):
    torch.manual_seed(0)
    logits = torch.randn(300, 77, device=dev) * 3
    mask = torch.rand(300, 77, device=dev) > 0.4
    mask[:, 0] = True
    lp, a = R.decode_step(logits, mask, "greedy", tanh_clipping=10.0, store_all_logp=True)
    ref = port.process_logits(logits.cpu(), mask.cpu(), 1.0, 10.0)
    torch.testing.assert_close(lp.cpu(), ref, rtol=1e-5, atol=1e-5)
    assert torch.equal(a.cpu(), ref.argmax(-1))
    lp2, a2 = R.decode_step(logits, mask, "sampling", temperature=2.0)
    assert mask.gather(1, a2[:, None]).all()
    ref2 = port.process_logits(logits.cpu(), mask.cpu(), 2.0, 0.0)
    torch.testing.assert_close(lp2.cpu(), ref2.gather(1, a2.cpu()[:, None]).squeeze(1), rtol=1e-5, atol=1e-5)


def test_reward_stats_and_global_mean(dev):
    torch.manual_seed(1)
    r = -torch.rand(100_003, device=dev) * 30
    s = R.reward_stats(r)
    assert s[2].item() == 100_003
    torch.testing.assert_close(s[0], r.double().sum(), rtol=1e-12, atol=1e-9)
    torch.testing.assert_close(s[1], (r.double() ** 2).sum(), rtol=1e-12, atol=1e-9)
    torch.testing.assert_close(R.global_reward_mean(r), r.double().mean(), rtol=1e-12, atol=1e-12)


def test_large_batch_greedy_properties(dev):
    """size-independent checks at a large batch: every tour is a permutation, reward equals
    the env's own replay of the returned tours, evaluate(ll) reproduces ll, and a sub-sample
    matches the oracle."""
    B, N = 8192, 100
    torch.manual_seed(1234)
    pol, env = load_policy("tsp", dev), R.TSPEnv(generator_params=dict(num_loc=N), device=dev)
    td = env.reset(batch_size=[B])
    with torch.inference_mode():
        out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
    _, r = env.replay(td, out["actions"], check=True)
    torch.testing.assert_close(r, out["reward"], rtol=1e-5, atol=1e-6)
    h = out["hidden"].node_embeddings
    ev = pol(td, env, actions=out["actions"], hidden=h)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=1e-6, atol=1e-6)
    idx = torch.arange(0, B, B // 32, device=dev)
    ref = port.am_rollout("tsp", port.tsp_reset(td["locs"][idx].cpu()), h[idx].cpu(), oracle_weights("tsp"),
                          actions=out["actions"][idx].cpu())
    torch.testing.assert_close(out["log_likelihood"][idx].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)


@pytest.mark.parametrize("N,B", ((2, 5), (3, 33), (16, 150), (17, 40), (32, 149), (33, 300), (64, 200), (65, 160),
                                 (80, 77), (97, 148), (112, 150)))
def test_tsp_greedy_compaction_kernel_sizes(N, B, dev):
    """The TSP greedy kernel (am_rollout_tsp.cu: 7 -> 4 -> 2 rows of node slots, compaction at 64 and 32 nodes left)
    at every stage boundary and at ragged batch sizes (fewer / more instances than SMs): valid tours, rewards equal to
    the environment's replay of them, log-likelihoods equal to an evaluate-mode pass and to the oracle on its own
    tours; tours equal to the generic register-file kernel's except for near-tie flips."""
    from rl4co_slap_b200 import _lib
    torch.manual_seed(100 + N)
    pol, env = load_policy("tsp", dev), R.TSPEnv(generator_params=dict(num_loc=N), device=dev)
    td = env.reset(batch_size=[B])
    with torch.inference_mode():
        out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
        h = out["hidden"].node_embeddings
        _lib.call("rl4co_set_rollout_impl", 2)
        try:
            gen = pol(td, env, decode_type="greedy", return_actions=True, hidden=h)
        finally:
            _lib.call("rl4co_set_rollout_impl", 0)
        ev = pol(td, env, actions=out["actions"], hidden=h)
    assert out["actions"].shape == (B, N)
    assert torch.equal(out["actions"].sort(1).values, torch.arange(N, device=dev).expand(B, N)), "not a permutation"
    _, r = env.replay(td, out["actions"], check=True)
    torch.testing.assert_close(r, out["reward"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=1e-5, atol=2e-5)
    same = (out["actions"] == gen["actions"]).all(1)
    # compaction kernel vs generic kernel: every diverging tour must split at a verified near tie of the oracle's
    # log-probs (a stage-boundary bug would split at an arbitrary step with a clear winner)
    n_div, gap = assert_divergences_are_near_ties("tsp", port.tsp_reset(td["locs"].cpu()), h, oracle_weights("tsp"),
                                                  gen["actions"], out["actions"])
    assert n_div == int((~same).sum())
    assert same.float().mean() >= 0.9, f"{n_div} of {B} tours flip at near ties (largest gap {gap:.2e})"
    torch.testing.assert_close(out["log_likelihood"][same], gen["log_likelihood"][same], rtol=1e-5, atol=2e-5)
    idx = torch.arange(0, B, max(1, B // 16), device=dev)
    ref = port.am_rollout("tsp", port.tsp_reset(td["locs"][idx].cpu()), h[idx].cpu(), oracle_weights("tsp"),
                          actions=out["actions"][idx].cpu())
    torch.testing.assert_close(out["log_likelihood"][idx].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)
    torch.testing.assert_close(out["reward"][idx].cpu(), ref["reward"], rtol=RTOL, atol=1e-6)
    # the same kernel in sampling mode (Philox inverse-CDF in slot order): feasible tours, ll of the sampled tour
    with torch.inference_mode():
        smp = pol(td, env, decode_type="sampling", return_actions=True, hidden=h)
        ev = pol(td, env, actions=smp["actions"], hidden=h)
    assert torch.equal(smp["actions"].sort(1).values, torch.arange(N, device=dev).expand(B, N)), "not a permutation"
    _, r = env.replay(td, smp["actions"], check=True)
    torch.testing.assert_close(r, smp["reward"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ev["log_likelihood"], smp["log_likelihood"], rtol=1e-5, atol=2e-5)
    if N > 3:
        assert not torch.equal(smp["actions"], out["actions"]), "sampling reproduced the greedy tours"


@pytest.mark.parametrize("M,nt", ((1, 1), (100, 3), (128 * 7 + 5, 5), (148 * 128 * 3 + 17, 4)))
def test_tensor_core_projection_matches_fp64(M, nt, dev):
    """K3 on tcgen05 (3xTF32 split, TMEM accumulation) vs an fp64 matmul and the fp32 SIMT kernel:
    ragged last tile, several tiles per CTA, every table count the policies use."""
    from rl4co_slap_b200 import _lib
    torch.manual_seed(M + nt)
    h = (torch.randn(M, 128, device=dev) * 3).contiguous()
    W = (torch.randn(nt * 128, 128, device=dev) * 0.1).contiguous()
    s = _lib.stream_ptr(dev)
    image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(nt)), device=dev)
    _lib.call("rl4co_am_pack_weights", nt, _lib.ptr(W), _lib.ptr(image), s)
    out_tc = torch.full((nt, M, 128), float("nan"), device=dev)
    out_simt = torch.empty(nt, M, 128, device=dev)
    _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(h), _lib.ptr(image), _lib.ptr(out_tc), s)
    _lib.call("rl4co_am_project", M, nt, _lib.ptr(h), _lib.ptr(W), _lib.ptr(out_simt), s)
    ref = (h.double() @ W.double().t()).view(M, nt, 128).permute(1, 0, 2)
    bound = 4e-6 * (h.double().abs() @ W.double().abs().t()).view(M, nt, 128).permute(1, 0, 2) + 1e-7
    assert not torch.isnan(out_tc).any()
    assert ((out_tc.double() - ref).abs() <= bound).all()          # fp32-level: ~1e-6 of sum |a||w|
    assert ((out_simt.double() - ref).abs() <= bound).all()


def test_policy_projection_paths_agree(dev):
    """the rollout through the tensor-core tables and through the fp32 SIMT tables: same tours, ll within 1e-5"""
    fx = fixture("tsp100")
    pol, env = load_policy("tsp", dev), R.TSPEnv(device=dev)
    td = env.reset(gen_td("tsp", fx, dev))
    h = t(fx["hidden"], dev)
    a = pol(td, env, decode_type="greedy", return_actions=True, hidden=h)
    pol.projection = "simt"
    b = pol(td, env, decode_type="greedy", return_actions=True, hidden=h)
    assert torch.equal(a["actions"], b["actions"])
    torch.testing.assert_close(a["log_likelihood"], b["log_likelihood"], rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_encoder_tensor_core_linears_match_pytorch(name, dev):
    """the encoder's K = 128 linears through rl4co_linear128_tc (bias / ReLU epilogue) vs torch fp32 F.linear"""
    torch.manual_seed(3)
    pol = load_policy(name, dev)
    g = R.TSPGenerator(num_loc=50) if name == "tsp" else R.CVRPGenerator(num_loc=50)
    td = R.get_env(name, device=dev).reset(g(300).to(dev))
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with torch.inference_mode():
            h_tc = pol.encoder(td)[0]
            for m in pol.encoder.modules():
                if isinstance(m, torch.nn.Linear):
                    m.tensor_core = False
                if hasattr(m, "Wqkv"):
                    m.fused_attention = False
            h_pt = pol.encoder(td)[0]
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    scale = h_pt.abs().max().item()
    assert (h_tc - h_pt).abs().max().item() <= 2e-5 * scale, ((h_tc - h_pt).abs().max().item(), scale)
    # and the fixture's reference embeddings (CPU fp32 reference encoder) through the tensor-core path
    fx = fixture(f"{name}50")
    pol2 = load_policy(name, dev)
    td2 = R.get_env(name, device=dev).reset(gen_td(name, fx, dev))
    with torch.inference_mode():
        h2 = pol2.encoder(td2)[0]
    ref = t(fx["hidden"], dev)
    assert (h2 - ref).abs().max().item() <= 2e-5 * ref.abs().max().item()


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_encoder_fused_skip_batchnorm_epilogue(name, dev):
    """eval-mode encoder layer: skip connection + batch norm (running statistics) fused into the epilogue of
    out_proj / the last feed-forward linear (rl4co_linear_tc_ex) vs the unfused PyTorch modules"""
    torch.manual_seed(11)
    pol = load_policy(name, dev).eval()
    for m in pol.encoder.modules():                 # non-trivial running statistics and affine
        if isinstance(m, torch.nn.BatchNorm1d):
            m.running_mean.copy_(torch.randn_like(m.running_mean) * 0.3)
            m.running_var.copy_(torch.rand_like(m.running_var) + 0.5)
            m.weight.data.copy_(torch.rand_like(m.weight) + 0.5)
            m.bias.data.copy_(torch.randn_like(m.bias) * 0.2)
    g = R.TSPGenerator(num_loc=50) if name == "tsp" else R.CVRPGenerator(num_loc=50)
    td = R.get_env(name, device=dev).reset(g(777).to(dev))       # 777 * 50 rows: a ragged last row tile
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with torch.inference_mode():
            h_fused = pol.encoder(td)[0]
            for m in pol.encoder.modules():
                if isinstance(m, torch.nn.Linear):
                    m.tensor_core = False
                if hasattr(m, "Wqkv"):
                    m.fused_attention = False
                if isinstance(m, torch.nn.Sequential) and len(m) == 4:
                    m.fused_epilogue = False
            h_pt = pol.encoder(td)[0]
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    scale = h_pt.abs().max().item()
    assert (h_fused - h_pt).abs().max().item() <= 2e-5 * scale, ((h_fused - h_pt).abs().max().item(), scale)


@pytest.mark.parametrize("B,N,amp", ((3, 100, 1.0), (7, 5, 1.0), (300, 50, 3.0), (150, 128, 1.0), (500, 101, 2.0), (1, 1, 1.0)))
def test_encoder_attention_kernels_match_fp64(B, N, amp, dev):
    """rl4co_encoder_attention_tc (the three kernels: Q K^T / Q K^T + P V on tcgen05, 3xTF32 / P V as 3xFP16) and rl4co_encoder_attention (SIMT) vs an fp64 softmax
    attention on the fused Wqkv layout [B, N, 3, 8, 16] (nn/attention.py:81-126, no mask)"""
    from rl4co_slap_b200 import _lib
    torch.manual_seed(B * 131 + N)
    qkv = torch.randn(B, N, 384, device=dev) * amp
    q, k, v = qkv.double().view(B, N, 3, 8, 16).permute(2, 0, 3, 1, 4).unbind(0)
    ref = (torch.softmax(q @ k.transpose(-1, -2) * 0.25, -1) @ v).transpose(1, 2).reshape(B, N, 128)
    variants = [("rl4co_encoder_attention_tc", 1), ("rl4co_encoder_attention", 0)]
    if N <= 112:
        variants.append(("rl4co_encoder_attention_tc", 2))     # P V on the tensor cores too (P in TMEM)
        variants.append(("rl4co_encoder_attention_tc", 3))     # ... with P and V as fp16 hi / lo pairs
        variants.append(("rl4co_encoder_attention_tc", 4))     # ... and four consumer groups (the default)
    try:
        for fn, impl in variants:
            _lib.call("rl4co_set_attention_impl", impl)
            out = torch.full((B, N, 128), float("nan"), device=dev)
            _lib.call(fn, B, N, _lib.ptr(qkv), _lib.ptr(out), _lib.stream_ptr(dev))
            err = (out.double() - ref).abs().max().item()
            assert err <= 1e-5 * max(1.0, amp ** 3) * max(1.0, ref.abs().max().item()), (fn, impl, err)
        if N > 112:
            _lib.call("rl4co_set_attention_impl", 2)
            with pytest.raises(RuntimeError, match="N <= 112"):
                _lib.call("rl4co_encoder_attention_tc", B, N, _lib.ptr(qkv), _lib.ptr(out), _lib.stream_ptr(dev))
    finally:
        _lib.call("rl4co_set_attention_impl", 0)


@pytest.mark.parametrize("M,n_out,n_in,relu", ((257, 128, 512, False), (1000, 512, 128, True), (130, 256, 256, True)))
def test_linear_tc_k_chunks_bias_relu(M, n_out, n_in, relu, dev):
    """rl4co_linear_tc: 128-wide K chunks accumulated in TMEM, bias / ReLU in the single epilogue"""
    from rl4co_slap_b200 import _lib
    torch.manual_seed(n_out + n_in)
    x = torch.randn(M, n_in, device=dev)
    W = torch.randn(n_out, n_in, device=dev) * 0.1
    b = torch.randn(n_out, device=dev)
    s = _lib.stream_ptr(dev)
    image = torch.empty((n_in // 128) * int(_lib.load().rl4co_am_weight_image_floats(n_out // 128)), device=dev)
    _lib.call("rl4co_linear_pack", n_out, n_in, _lib.ptr(W), _lib.ptr(image), s)
    y = torch.full((M, n_out), float("nan"), device=dev)
    _lib.call("rl4co_linear_tc", M, n_out, n_in, _lib.ptr(x), _lib.ptr(image), _lib.ptr(b), int(relu), _lib.ptr(y), s)
    ref = x.double() @ W.double().t() + b.double()
    if relu:
        ref = ref.clamp_min(0)
    bound = 4e-6 * (x.double().abs() @ W.double().abs().t() + b.double().abs()) + 1e-6
    assert ((y.double() - ref).abs() <= bound).all()


@pytest.mark.parametrize("M,n_in,bias", ((1, 2, True), (1000, 3, True), (16384 * 100 + 5, 2, True), (777, 4, False), (50, 8, True)))
def test_linear_narrow_init_embedding(M, n_in, bias, dev):
    """rl4co_linear_narrow (the encoder's init embedding, 2..8 features -> 128) against fp64 and torch's F.linear"""
    from rl4co_slap_b200 import _lib
    torch.manual_seed(M % 97 + n_in)
    x = torch.rand(M, n_in, device=dev) * 3 - 1
    lin = torch.nn.Linear(n_in, 128, bias=bias).to(dev)
    y = torch.full((M, 128), float("nan"), device=dev)
    _lib.call("rl4co_linear_narrow", M, n_in, _lib.ptr(x), _lib.ptr(lin.weight.detach()),
              _lib.ptr(lin.bias.detach()) if bias else None, _lib.ptr(y), _lib.stream_ptr(dev))
    ref = x.double() @ lin.weight.double().t() + (lin.bias.double() if bias else 0)
    mag = x.double().abs() @ lin.weight.double().abs().t() + (lin.bias.double().abs() if bias else 0)
    assert ((y.double() - ref).abs() <= 2.5e-7 * mag + 1e-9).all()
    with torch.inference_mode():
        torch.testing.assert_close(R.encoder.narrow_linear(x, lin), y, rtol=0, atol=0)        # the module path is this kernel
        torch.testing.assert_close(y, torch.nn.functional.linear(x, lin.weight, lin.bias), rtol=1e-6, atol=1e-6)


def _chi2(counts, probs, n):
    keep = probs * n > 5
    return float((((counts - n * probs) ** 2) / (n * probs).clamp_min(1e-9))[keep].sum()), int(keep.sum())


def test_cvrp_sampling_distribution_against_oracle_probabilities(dev):
    """BASELINE configs[3] is CVRP *sampling*: per-step action frequencies of the kernel's Gumbel-max draws against
    the oracle's probabilities (utils/decoding.py:383-397) - step 0 (everyone at the depot), step 1 conditioned on the
    most frequent first customer, and for POMO multistart the first SAMPLED step of one forced start; mean reward within
    4 standard errors of the oracle's own multinomial sampling."""
    fx = fixture("cvrp20")
    pol, env = load_policy("cvrp", dev), R.CVRPEnv(device=dev)
    W = oracle_weights("cvrp")
    reps = 8192
    g1 = {k: t(fx[f"gen.{k}"])[:1] for k in ("locs", "depot", "demand")}
    td = env.reset(R.TensorDict({k: v.expand(reps, *v.shape[1:]).contiguous().to(dev) for k, v in g1.items()},
                                batch_size=[reps]))
    h1 = t(fx["hidden"])[:1]
    hid = h1.expand(reps, -1, -1).contiguous().to(dev)
    torch.manual_seed(17)
    out = pol(td, env, decode_type="sampling", return_actions=True, hidden=hid)
    env.check_solution_validity(td, out["actions"])
    ev = pol(td, env, actions=out["actions"], hidden=hid)
    torch.testing.assert_close(ev["log_likelihood"], out["log_likelihood"], rtol=1e-6, atol=2e-6)
    N = h1.shape[1]
    acts = out["actions"].cpu()
    st1 = port.cvrp_reset(g1["depot"], g1["locs"], g1["demand"], 1.0)
    # step 0 and step 1 | most frequent first action: oracle log-probs along that prefix
    a0 = int(torch.bincount(acts[:, 0], minlength=N).argmax())
    prefix = torch.tensor([[a0, 0]])
    ref = port.am_rollout("cvrp", st1, h1, W, actions=prefix, return_all_logp=True, check_solution=False, max_steps=1)
    p0, p1 = ref["logprobs"][0, 0].exp().double(), ref["logprobs"][0, 1].exp().double()
    c0 = torch.bincount(acts[:, 0], minlength=N).double()
    chi, dof = _chi2(c0, p0, reps)
    assert chi < 40 + 3 * dof, (chi, dof)
    grp = acts[acts[:, 0] == a0]
    c1 = torch.bincount(grp[:, 1], minlength=N).double()
    chi, dof = _chi2(c1, p1, grp.shape[0])
    assert chi < 40 + 3 * dof, (chi, dof, grp.shape[0])
    assert c1[p1 == 0].sum() == 0, "a masked node was sampled"
    # mean reward against the oracle's own torch.multinomial sampling of the same instance
    torch.manual_seed(5)
    n_ref = 1024
    stn = port.cvrp_reset(g1["depot"].expand(n_ref, -1).contiguous(), g1["locs"].expand(n_ref, -1, -1).contiguous(),
                          g1["demand"].expand(n_ref, -1).contiguous(), 1.0)
    refs = port.am_rollout("cvrp", stn, h1.expand(n_ref, -1, -1), W, "sampling")
    a, b = out["reward"].double().cpu(), refs["reward"].double()
    se = (a.var() / a.numel() + b.var() / b.numel()).sqrt().item()
    assert abs(a.mean().item() - b.mean().item()) < 4 * se, (a.mean().item(), b.mean().item(), se)
    # POMO multistart sampling: start s is forced to customer s + 1 with log-prob 0; the next action is the first draw
    S, repm = 4, 2048
    tdm = env.reset(R.TensorDict({k: v.expand(repm, *v.shape[1:]).contiguous().to(dev) for k, v in g1.items()},
                                 batch_size=[repm]))
    outm = pol(tdm, env, decode_type="multistart_sampling", num_starts=S, return_actions=True, hidden=hid[:repm])
    am = outm["actions"].cpu().view(S, repm, -1)
    for s in range(S):
        assert (am[s, :, 0] == s + 1).all()
        refm = port.am_rollout("cvrp", st1, h1, W, actions=torch.tensor([[s + 1, 0]]), return_all_logp=True,
                               check_solution=False, max_steps=1)
        pm = refm["logprobs"][0, 1].exp().double()
        chi, dof = _chi2(torch.bincount(am[s, :, 1], minlength=N).double(), pm, repm)
        assert chi < 40 + 3 * dof, (s, chi, dof)


def test_slap_large_batch_parity_on_a_subsample(dev):
    """BASELINE configs[4] is a large-batch SLAP sweep: B = 16 384 instances generated on the device (every instance
    has its own picklist), greedy and sampled rollouts; a sub-sample of 64 against the oracle - greedy assignment bit
    exact, evaluate-mode ll / reward of the sampled assignments within 1e-5 - and the reward of EVERY rollout against the
    environment's own get_reward."""
    B, P = 16384, 20
    torch.manual_seed(77)
    env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
    pol = load_policy("slap", dev)
    gen = env.generator(B)
    td = env.reset(gen.to(dev))
    with torch.inference_mode():
        out = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
        h = out["hidden"].node_embeddings
        smp = pol(td, env, decode_type="sampling", return_actions=True, hidden=h)
    assert out["actions"].shape == (B, P)
    for o in (out, smp):
        srt = o["actions"].sort(1).values
        assert (srt[:, 1:] != srt[:, :-1]).all() and (o["actions"] > 0).all(), "a location was assigned twice / the depot"
        assert torch.equal(o["assignment"].long(), o["actions"]), "product t is assigned at step t"
        _, r = env.replay(td, o["actions"])
        torch.testing.assert_close(r, o["reward"], rtol=1e-5, atol=1e-5)
    assert smp["actions"].unique(dim=0).shape[0] > B // 2
    idx = torch.arange(0, B, B // 64, device=dev)
    cpu = {k: gen[k][idx].cpu() for k in ("locs", "freq", "assignment", "picklist", "depot_loc_dist")}

    def state():
        return port.slap_reset(cpu["locs"], cpu["assignment"].clone(), P,
                               {k: cpu[k] for k in ("freq", "picklist", "depot_loc_dist")})

    W = oracle_weights("slap")
    ref = port.am_rollout("slap", state(), h[idx].cpu(), W, "greedy")
    assert torch.equal(out["actions"][idx].cpu(), ref["actions"])
    torch.testing.assert_close(out["reward"][idx].cpu(), ref["reward"], rtol=RTOL, atol=1e-5)
    torch.testing.assert_close(out["log_likelihood"][idx].cpu(), ref["log_likelihood"], rtol=RTOL, atol=2e-5)
    refs = port.am_rollout("slap", state(), h[idx].cpu(), W, actions=smp["actions"][idx].cpu())
    torch.testing.assert_close(smp["reward"][idx].cpu(), refs["reward"], rtol=RTOL, atol=1e-5)
    torch.testing.assert_close(smp["log_likelihood"][idx].cpu(), refs["log_likelihood"], rtol=RTOL, atol=2e-5)


def test_train_mode_batch_norm_sees_the_whole_batch_once(dev):
    """chunked rollouts must not chunk a train-mode batch-norm encoder: same hidden states (hence tours) as an unchunked
    call, and the running statistics move exactly once per forward"""
    torch.manual_seed(21)
    env = R.TSPEnv(generator_params=dict(num_loc=20), device=dev)
    td = env.reset(batch_size=[96])
    pol = R.AttentionModelPolicy(env_name="tsp", num_encoder_layers=2).to(dev).train()
    bn = next(m for m in pol.encoder.modules() if isinstance(m, torch.nn.BatchNorm1d))
    sd = {k: v.clone() for k, v in pol.state_dict().items()}
    with torch.no_grad():
        whole = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
        n1 = int(bn.num_batches_tracked)
        pol.load_state_dict(sd)
        pol.rollout_chunk = 32
        parts = pol(td, env, decode_type="greedy", return_actions=True, return_hidden=True)
        n2 = int(bn.num_batches_tracked)
    assert n1 == n2 == int(sd[[k for k in sd if k.endswith("num_batches_tracked")][0]]) + 1
    torch.testing.assert_close(parts["hidden"].node_embeddings, whole["hidden"].node_embeddings, rtol=0, atol=0)
    assert torch.equal(parts["actions"], whole["actions"])


@pytest.mark.parametrize("name,num_loc,B", (("tsp", 50, 512), ("slap", None, 64)))
def test_graphed_rollout_equals_eager_and_replays_on_new_data(name, num_loc, B, dev):
    """policy.graphed(): encoder + projection + rollout captured as one CUDA graph; replays on fresh instances give exactly
    the eager results, the captured output tensors are reused, and device-side assertions still surface"""
    torch.manual_seed(31)
    gp = dict(num_loc=num_loc) if name == "tsp" else {"n_aisles": 10, "n_locs": 10}
    env = R.get_env(name, device=dev, generator_params=gp, check_solution=False)
    pol = load_policy(name, dev)
    td0 = env.reset(env.generator(B).to(dev))
    with torch.inference_mode():
        fast = pol.graphed(td0, env, decode_type="greedy", return_actions=True)
        for seed in (1, 2):
            torch.manual_seed(seed)
            td = env.reset(env.generator(B).to(dev))
            want = pol(td, env, decode_type="greedy", return_actions=True)
            got = fast(td)
            assert torch.equal(got["actions"], want["actions"])
            assert torch.equal(got["reward"], want["reward"]) and torch.equal(got["log_likelihood"], want["log_likelihood"])
        first_ptr = got["actions"].data_ptr()
        assert fast(td)["actions"].data_ptr() == first_ptr
    with pytest.raises(NotImplementedError):
        pol.graphed(td0, env, decode_type="sampling")
    if name == "tsp":   # a NaN coordinate -> NaN logits: the status word is read after the replay
        bad = td.clone()
        bad["locs"] = td["locs"].clone()
        bad["locs"][3, 7, 0] = float("nan")
        with pytest.raises(AssertionError, match="Logits contain NaNs|infeasible action selected"):
            with torch.inference_mode():
                fast(bad)
        with torch.inference_mode():
            again = fast(td)                    # and the next replay is clean again
        assert torch.equal(again["actions"], want["actions"])
