"""Randomised soak of the round-2 kernels (not collected by pytest; run on a GPU box):
    python tools/soak_round2.py [seconds per part, default 60]
  k3   - rl4co_am_project_tc / rl4co_linear_tc (FP16 form): random row counts (ragged tiles), 1..8 tables, input magnitudes from
         1e-6 to 1e5 incl. tiles of zeros and single huge outliers, against fp64;
  slap - am_rollout_slap.cu against the generic register-file kernel on random warehouse shapes (N = 6..128, P = 2..32), random
         per-instance hidden states at random scales (collapsed attention included), greedy / sampling / evaluate: evaluate
         log-likelihoods of the same assignments must agree to the fp32 class of the inputs, rewards to 1e-5.
  linear - rl4co_linear_tc_ex: random widths (K = 128 on FP16 halves, K chunks on TF32), bias / ReLU / skip row / affine, against fp64;
  multistart - am_rollout_ms_tc.cu against the SIMT kernel: random tsp / cvrp sizes, batch and start counts, greedy tours equal
         up to near ties with equal log-likelihoods and rewards, sampled tours valid.  SOAK_MS_SCALE=1 also scales the encoder
         output by 10^-2.5..10^0.5 (the kernel's fp16 hi / lo operands carry no range scale): 4 758 cases, the log-likelihoods
         of equal tours always within tolerance; 2 cases at scales < 0.01 fell below the 90 % equal-tours line, where the logits
         are uniform to 1e-6 and every step is a near tie.
  backward - am_backward.cu against autograd through the batched PyTorch pass (log-likelihoods, gradients of the encoder output
         and of every decoder weight), random tsp / cvrp sizes, single start and multistart;
  start nodes - multistart with caller-given start nodes (sdvrp, op, cvrp, tsp): every start must be the evaluate-mode episode of
         its instance with that first action (rewards, per-step log-probs).
  consistency - all seven environments at random sizes: sampled and greedy rollouts pass the environment's validity check, the
         fused replay reproduces their reward and the evaluate path their log-likelihood.
  attention - rl4co_encoder_attention_tc (default kernel: both GEMMs on tcgen05, P V as 3 x fp16) against fp64 softmax attention:
         random batch / node counts (1..112 on that kernel, above it the scores-only kernel), input magnitudes 1e-3..8, rows of
         zeros, one dominant key.
  firstctx - tsp with four tables (the kernels form the first node's context row themselves) against five tables: random sizes,
         greedy tours equal or split at a log-prob gap < 1e-4, evaluate log-likelihoods to 1e-5, on all three kernels.
Prints one line per case that exceeds its bound and a summary; exit code 1 if any did."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib

dev = torch.device("cuda:0")
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
only = set(sys.argv[2:])                # optional part names: k3 slap linear multistart backward starts consistency attention firstctx
bad = 0
g = torch.Generator(device="cpu").manual_seed(20261105)
rnd = lambda lo, hi: int(torch.randint(lo, hi + 1, (1,), generator=g))

# ---------------------------------------------------------------- K3 / linear, FP16 form
t0, n = time.time(), 0
st = _lib.stream_ptr(dev)
while (not only or "k3" in only) and time.time() - t0 < budget:
    M, nt = rnd(1, 70000), rnd(1, 8)
    mag = 10.0 ** (rnd(-60, 50) / 10.0)
    x = torch.randn(M, 128, device=dev) * mag
    kind = rnd(0, 3)
    if kind == 1 and M > 300:
        x[128:256] = 0                                          # a whole tile of zeros
    if kind == 2:
        x[rnd(0, M - 1), rnd(0, 127)] = mag * 3e4               # one outlier sets the tile's scale
    if kind == 3:
        x *= torch.logspace(-6, 0, 128, device=dev)             # columns spanning six decades
    W = torch.randn(nt * 128, 128, device=dev) * 10.0 ** (rnd(-30, 10) / 10.0)
    image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(nt)), device=dev)
    _lib.call("rl4co_am_pack_weights", nt, _lib.ptr(W), _lib.ptr(image), st)
    out = torch.full((nt, M, 128), float("nan"), device=dev)
    _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(x), _lib.ptr(image), _lib.ptr(out), st)
    rows = torch.randperm(M, device=dev)[:2048]
    ref = (x[rows].double() @ W.double().t()).view(-1, nt, 128).permute(1, 0, 2)
    magn = (x[rows].double().abs() @ W.double().abs().t()).view(-1, nt, 128).permute(1, 0, 2)
    # 3-term split: <= 8.3e-7 per product (measured over all mantissas); the 24 MMAs of a tile accumulate into the fp32
    # TMEM accumulator one after the other, each rounding at 2^-23..-24 of the running sum - up to ~2e-6 of a row that one
    # large product dominates (measured); plus the fp16 floor of entries far below the largest of their tile of x / their
    # table of W: half an fp16-subnormal ulp, 2^-25 of a scaled value whose largest is >= 2^13, on either operand
    amax = x.abs().max().double() * W.abs().max().double()
    err = (out[:, rows].double() - ref).abs()
    bound = 3.5e-6 * magn + 2 * 128 * 2.0 ** -25 * 2.0 ** -13 * amax * 1.5 + 1e-30
    if not bool((err <= bound).all()) or not bool(torch.isfinite(out).all()):
        bad += 1
        print(f"K3 FAIL M={M} nt={nt} mag={mag:.2e} kind={kind}: worst excess {float((err / bound).max()):.2f}x", flush=True)
    n += 1
print(f"k3: {n} random cases, {bad} failures", flush=True)

# ---------------------------------------------------------------- SLAP kernel vs generic kernel
t0, n, bad_k3 = time.time(), 0, bad
with torch.inference_mode():
    while (not only or "slap" in only) and time.time() - t0 < budget:
        na, nl = rnd(2, 16), rnd(3, 8)
        N = na * nl
        P = rnd(2, min(32, N - 1))
        O, K = rnd(1, 30), rnd(1, 8)
        B = rnd(1, 900)
        env = R.SLAPEnv(generator_params={"n_aisles": na, "n_locs": nl, "n_products": P, "max_orders": O,
                                          "max_products_in_order": K}, device=dev)
        pol = R.AttentionModelPolicy(env_name="slap", num_encoder_layers=1, feedforward_hidden=128,
                                     use_graph_context=bool(rnd(0, 3))).to(dev).eval()
        td = env.reset(env.generator(B).to(dev))
        h = pol.encoder(td)[0].float()
        scale = 10.0 ** (rnd(-10, 14) / 10.0)
        h = ((h + 0.3 * h.std() * torch.randn_like(h)) * scale).contiguous()
        res = {}
        for impl in (0, 2):
            _lib.call("rl4co_set_rollout_impl", impl)
            torch.manual_seed(n)
            smp = pol(td, env, decode_type="sampling", hidden=h, return_actions=True)
            if impl == 0:
                acts = smp["actions"]
            res[impl] = (smp, pol(td, env, actions=acts, hidden=h), pol(td, env, decode_type="greedy", hidden=h, return_actions=True))
        _lib.call("rl4co_set_rollout_impl", 0)
        a, b = res[0][1], res[2][1]
        ll_a, ll_b = a["log_likelihood"].double(), b["log_likelihood"].double()
        # both kernels are fp32 implementations of an ill-conditioned map when attention collapses: scores of size s carry
        # eps * s of rounding, which the exponential turns into a relative error of the weights
        smax = float(h.abs().max()) ** 2
        tol = 1e-5 * ll_b.abs() + 3e-5 + P * 6e-8 * smax * 8
        ok = bool(((ll_a - ll_b).abs() <= tol).all()) and bool(torch.allclose(a["reward"], b["reward"], rtol=1e-5, atol=1e-5))
        same_g = (res[0][2]["actions"] == res[2][2]["actions"]).all(1).float().mean().item()
        same_s = (res[0][0]["actions"] == res[2][0]["actions"]).all(1).float().mean().item()
        srt = res[0][0]["actions"].sort(1).values
        valid = bool((srt[:, 1:] != srt[:, :-1]).all()) and bool((res[0][0]["actions"] > 0).all()) and \
            bool((res[0][0]["actions"] < N).all())
        if not (ok and valid):
            bad += 1
            print(f"SLAP FAIL N={N} P={P} O={O} K={K} B={B} scale={scale:.2f}: ll diff {float((ll_a - ll_b).abs().max()):.3e} "
                  f"(tol {float(tol.min()):.1e}), valid={valid}, greedy same {same_g:.3f}, sampled same {same_s:.3f}", flush=True)
        elif same_g < 0.9 and scale < 3:
            print(f"slap note N={N} P={P} B={B} scale={scale:.2f}: greedy same {same_g:.3f}, sampled same {same_s:.3f}", flush=True)
        n += 1
print(f"slap: {n} random cases, {bad - bad_k3} failures", flush=True)

# ---------------------------------------------------------------- linear layers with epilogues (FP16 K = 128, TF32 K chunks)
t0, n, bad_prev = time.time(), 0, bad
while (not only or "linear" in only) and time.time() - t0 < budget:
    M = rnd(1, 40000)
    n_out, n_in = 128 * rnd(1, 4), 128 * rnd(1, 4)
    relu, use_res, use_bn, use_bias = rnd(0, 1), rnd(0, 1), rnd(0, 1), rnd(0, 1)
    mag = 10.0 ** (rnd(-30, 30) / 10.0)
    x = torch.randn(M, n_in, device=dev) * mag
    Wl = torch.randn(n_out, n_in, device=dev) * 0.1
    bias = torch.randn(n_out, device=dev) * mag if use_bias else None
    resid = torch.randn(M, n_out, device=dev) * mag if use_res else None
    sc_, sh_ = (torch.rand(n_out, device=dev) + 0.5, torch.randn(n_out, device=dev) * mag) if use_bn else (None, None)
    image = torch.empty((n_in // 128) * int(_lib.load().rl4co_am_weight_image_floats(n_out // 128)), device=dev)
    _lib.call("rl4co_linear_pack", n_out, n_in, _lib.ptr(Wl), _lib.ptr(image), st)
    y = torch.full((M, n_out), float("nan"), device=dev)
    _lib.call("rl4co_linear_tc_ex", M, n_out, n_in, _lib.ptr(x), _lib.ptr(image), _lib.ptr(bias), relu, _lib.ptr(resid),
              _lib.ptr(sc_), _lib.ptr(sh_), _lib.ptr(y), st)
    rows = torch.randperm(M, device=dev)[:1024]
    pre = x[rows].double() @ Wl.double().t() + (bias.double() if use_bias else 0)
    magn = x[rows].double().abs() @ Wl.double().abs().t() + (bias.double().abs() if use_bias else 0)
    ref = pre.clamp_min(0) if relu else pre
    if use_res:
        ref, magn = ref + resid[rows].double(), magn + resid[rows].double().abs()
    if use_bn:
        ref, magn = ref * sc_.double() + sh_.double(), magn * sc_.double() + sh_.double().abs()
    err = (y[rows].double() - ref).abs()
    bound = (4e-6 if n_in == 128 else 8e-6) * magn + 1e-30     # (TF32 K chunks: 48 accumulations per chunk)
    if not bool((err <= bound).all()) or not bool(torch.isfinite(y).all()):
        bad += 1
        print(f"LINEAR FAIL M={M} {n_in}->{n_out} relu={relu} res={use_res} bn={use_bn} bias={use_bias} mag={mag:.1e}: "
              f"worst excess {float((err / bound).max()):.2f}x", flush=True)
    n += 1
print(f"linear: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- multistart tensor-core kernel vs the SIMT kernel
t0, n, bad_prev = time.time(), 0, bad
with torch.inference_mode():
    while (not only or "multistart" in only) and time.time() - t0 < budget:
        name = ("tsp", "cvrp")[rnd(0, 1)]
        num_loc = rnd(5, 100)
        B, S = rnd(1, 160), rnd(1, min(128, num_loc))
        env = R.get_env(name, generator_params=dict(num_loc=num_loc), device=dev, check_solution=False)
        pol = R.POMOPolicy(env_name=name, num_encoder_layers=1).to(dev).eval()
        td = env.reset(env.generator(B).to(dev))
        hs = 10.0 ** (rnd(-25, 5) / 10.0) if os.environ.get("SOAK_MS_SCALE") else 1.0   # embeddings at another magnitude
        h = (pol.encoder(td)[0].float() * hs).contiguous()
        out = {}
        for impl in (0, 2):
            _lib.call("rl4co_set_rollout_impl", impl)
            torch.manual_seed(n)
            out[impl] = (pol(td, env, decode_type="multistart_greedy", num_starts=S, hidden=h, return_actions=True),
                         pol(td, env, decode_type="multistart_sampling", num_starts=S, hidden=h, return_actions=True))
        _lib.call("rl4co_set_rollout_impl", 0)
        ok = True
        why = ""
        for k in (0, 1):
            a, b = out[0][k], out[2][k]
            T = max(a["actions"].shape[1], b["actions"].shape[1])
            pa = torch.nn.functional.pad(a["actions"], (0, T - a["actions"].shape[1]))
            pb = torch.nn.functional.pad(b["actions"], (0, T - b["actions"].shape[1]))
            same = (pa == pb).all(1)
            if k == 0 and float(same.float().mean()) < 0.9:
                ok = False
                why += f" equal greedy tours {float(same.float().mean()):.2f}"
            if bool(same.any()):
                if not torch.allclose(a["log_likelihood"][same], b["log_likelihood"][same], rtol=1e-5, atol=5e-5):
                    ok = False
                    why += f" ll diff {float((a['log_likelihood'][same] - b['log_likelihood'][same]).abs().max()):.2e}"
                ok &= bool(torch.allclose(a["reward"][same], b["reward"][same], rtol=1e-5, atol=1e-5))
            # the tensor-core kernel samples by inverse CDF, the SIMT kernel by Gumbel-max: different draws, both must be tours
            if name == "tsp":
                srt = a["actions"].sort(1).values
                ok &= bool((srt == torch.arange(srt.shape[1], device=dev)).all())
            else:
                env2 = R.get_env(name, generator_params=dict(num_loc=num_loc), device=dev, check_solution=True)
                env2.check_solution_validity(R.batchify(td, S), a["actions"])
        if not ok:
            bad += 1
            print(f"MS_TC FAIL {name}{num_loc} B={B} S={S} h scale {hs:.3g}:{why}", flush=True)
        n += 1
print(f"multistart: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- backward kernel vs autograd through the PyTorch pass
from rl4co_slap_b200.training import kernel_backward_supported, kernel_log_likelihood, teacher_forced_log_likelihood


def _grads(fn, pol, h, weights):
    h = h.detach().clone().requires_grad_(True)
    for p_ in pol.decoder.parameters():
        p_.grad = None
    ll = fn(h)
    (weights * ll).sum().backward()
    out = {"h": h.grad.detach().clone()}
    for name_, p_ in pol.decoder.named_parameters():
        if p_.requires_grad:
            out[name_] = None if p_.grad is None else p_.grad.detach().clone()
    return ll.detach(), out


t0, n, bad_prev = time.time(), 0, bad
while (not only or "backward" in only) and time.time() - t0 < budget:
    name = ("tsp", "cvrp")[rnd(0, 1)]
    num_loc = rnd(5, 100)
    S = 0 if rnd(0, 1) else rnd(2, min(12, num_loc))
    B = rnd(1, 24)
    pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=1, use_graph_context=bool(rnd(0, 2))).to(dev)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=num_loc))
    td = env.reset(env.generator(B).to(dev))
    with torch.no_grad():
        h0 = pol.encoder(td)[0].float()
        kw = dict(decode_type="multistart_sampling", num_starts=S) if S else dict(decode_type="sampling")
        out = pol(td, env, return_actions=True, hidden=h0, **kw)
    if not kernel_backward_supported(pol, name, h0):
        continue
    w_ = torch.randn(out["actions"].shape[0], device=dev)
    ll_ref, g_ref = _grads(lambda hh: teacher_forced_log_likelihood(pol, td, name, hh, out["actions"], num_starts=S), pol, h0, w_)
    ll_ker, g_ker = _grads(lambda hh: kernel_log_likelihood(pol, td, name, hh, out["actions"], num_starts=S), pol, h0, w_)
    ok = bool(torch.allclose(ll_ker, ll_ref, rtol=1e-5, atol=3e-5))
    worst = ""
    for key, want in g_ref.items():
        got = g_ker.get(key)
        if want is None or float(want.abs().max()) == 0.0:
            ok &= got is None or float(got.abs().max()) < 1e-5
            continue
        e_, s_ = float((got - want).abs().max()), float(want.abs().max())
        if e_ > 3e-4 * s_:
            ok, worst = False, f"{key}: {e_:.2e} of {s_:.2e}"
    if not ok:
        bad += 1
        print(f"BACKWARD FAIL {name}{num_loc} B={B} S={S} {worst}", flush=True)
    n += 1
print(f"backward: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- sdvrp multistart / given start nodes vs the evaluate path
t0, n, bad_prev = time.time(), 0, bad
with torch.inference_mode():
    while (not only or "starts" in only) and time.time() - t0 < budget:
        name = ("sdvrp", "op", "cvrp", "tsp")[rnd(0, 3)]
        num_loc = rnd(6, 60)
        B, S = rnd(1, 40), rnd(1, min(16, num_loc))
        env = R.get_env(name, device=dev, generator_params=dict(num_loc=num_loc), check_solution=False)
        pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=1).to(dev).eval()
        td = env.reset(env.generator(B).to(dev))
        h = pol.encoder(td)[0].float().contiguous()
        lo = 0 if name == "tsp" else 1
        if name == "op":                                    # the reference draws them from the feasible nodes
            feas = td["action_mask"][..., 1:].float()
            starts = (torch.multinomial(feas, S, replacement=True) + 1).t().reshape(-1)
        else:
            starts = torch.randint(lo, num_loc + lo, (S * B,), device=dev)
        fn = lambda td_, env_, k: starts
        out = pol(td, env, decode_type="multistart_greedy", num_starts=S, hidden=h, return_actions=True,
                  return_sum_log_likelihood=False, select_start_nodes_fn=fn)
        ok = bool((out["actions"][:, 0] == starts).all()) and bool((out["log_likelihood"][:, 0] == 0).all())
        ev = pol(R.batchify(td, S), env, actions=out["actions"], hidden=R.batchify(h, S), return_sum_log_likelihood=False)
        ok &= bool(torch.allclose(ev["reward"], out["reward"], rtol=1e-5, atol=1e-5))
        ok &= bool(torch.allclose(ev["log_likelihood"][:, 1:], out["log_likelihood"][:, 1:], rtol=1e-4, atol=5e-5))
        if not ok:
            bad += 1
            print(f"STARTS FAIL {name}{num_loc} B={B} S={S}: reward diff {float((ev['reward'] - out['reward']).abs().max()):.2e} "
                  f"logp diff {float((ev['log_likelihood'][:, 1:] - out['log_likelihood'][:, 1:]).abs().max()):.2e}", flush=True)
        n += 1
print(f"start nodes: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- every env: sampled rollouts are valid and re-evaluate to themselves
t0, n, bad_prev = time.time(), 0, bad
with torch.inference_mode():
    while (not only or "consistency" in only) and time.time() - t0 < budget:
        name = ("tsp", "cvrp", "sdvrp", "op", "pctsp", "spctsp", "slap")[rnd(0, 6)]
        if name == "slap":
            na, nl = rnd(2, 12), rnd(3, 10)
            num_loc = na * nl
            env = R.SLAPEnv(generator_params={"n_aisles": na, "n_locs": nl, "n_products": rnd(2, min(30, num_loc - 1))}, device=dev)
        else:
            num_loc = rnd(4, 120)
            env = R.get_env(name, device=dev, generator_params=dict(num_loc=num_loc))
        B = rnd(1, 300)
        pol = R.AttentionModelPolicy(env_name=name, num_encoder_layers=1).to(dev).eval()
        td = env.reset(env.generator(B).to(dev))
        h = pol.encoder(td)[0].float().contiguous()
        torch.manual_seed(n)
        smp = pol(td, env, decode_type="sampling", hidden=h, return_actions=True)
        grd = pol(td, env, decode_type="greedy", hidden=h, return_actions=True)
        ok = True
        try:
            for o in (smp, grd):
                if name == "slap":
                    srt = o["actions"].sort(1).values
                    assert bool((srt[:, 1:] != srt[:, :-1]).all()) and bool((o["actions"] > 0).all())
                else:
                    env.check_solution_validity(td, o["actions"])
                ev = pol(td, env, actions=o["actions"], hidden=h)
                assert torch.allclose(ev["reward"], o["reward"], rtol=1e-5, atol=1e-5), "reward"
                assert torch.allclose(ev["log_likelihood"], o["log_likelihood"], rtol=1e-4, atol=1e-4), \
                    f"ll {float((ev['log_likelihood'] - o['log_likelihood']).abs().max()):.2e}"
                _, r = env.replay(td, o["actions"])
                assert torch.allclose(r, o["reward"], rtol=1e-5, atol=1e-5), "replay reward"
        except AssertionError as exc:
            ok = False
            print(f"CONSISTENCY FAIL {name}{num_loc} B={B}: {str(exc)[:120]}", flush=True)
        if not ok:
            bad += 1
        n += 1
print(f"consistency: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- encoder attention vs fp64
t0, n, bad_prev = time.time(), 0, bad
while (not only or "attention" in only) and time.time() - t0 < budget:
    B, N = rnd(1, 600), rnd(1, 128)
    amp = 10.0 ** (rnd(-30, 9) / 10.0)
    qkv = torch.randn(B, N, 384, device=dev) * amp
    kind = rnd(0, 3)
    if kind == 1:
        qkv[rnd(0, B - 1), rnd(0, N - 1)] = 0                   # a node whose q, k, v are all zero
    if kind == 2:
        qkv[:, rnd(0, N - 1), 128:256] *= 4.0                   # one key dominates every row
    if kind == 3:
        qkv[..., 256:] *= 10.0 ** (rnd(-30, 20) / 10.0)         # values at another scale than queries / keys
    q, k, v = qkv.double().view(B, N, 3, 8, 16).permute(2, 0, 3, 1, 4).unbind(0)
    ref = (torch.softmax(q @ k.transpose(-1, -2) * 0.25, -1) @ v).transpose(1, 2).reshape(B, N, 128)
    out = torch.full((B, N, 128), float("nan"), device=dev)
    _lib.call("rl4co_encoder_attention_tc", B, N, _lib.ptr(qkv), _lib.ptr(out), st)
    # scores carry an absolute error of ~1e-6 |q||k| (3xTF32), i.e. a relative one on the weights: 1e-5 max(1, amp^2) of the
    # largest value a row can mix, plus the fp16 floor of the values (2^-25 of their magnitude class)
    vmax = float(v.abs().max())
    bound = 1e-5 * max(1.0, float(q.abs().max()) * float(k.abs().max()) / 16) * vmax + 1e-7 * vmax + 1e-30
    err = float((out.double() - ref).abs().max())
    if not (err <= bound) or not bool(torch.isfinite(out).all()):
        bad += 1
        print(f"ATTENTION FAIL B={B} N={N} amp={amp:.3g} kind={kind}: err {err:.3e} bound {bound:.3e}", flush=True)
    n += 1
print(f"attention: {n} random cases, {bad - bad_prev} failures", flush=True)

# ---------------------------------------------------------------- tsp: four tables + in-kernel first-node row vs five tables
t0, n, bad_prev = time.time(), 0, bad
with torch.inference_mode():
    while (not only or "firstctx" in only) and time.time() - t0 < budget:
        num_loc, B = rnd(2, 120), rnd(1, 400)
        env = R.get_env("tsp", device=dev, generator_params=dict(num_loc=num_loc))
        pol = R.AttentionModelPolicy(env_name="tsp", num_encoder_layers=1).to(dev).eval()
        td = env.reset(env.generator(B).to(dev))
        h = (pol.encoder(td)[0].float() * 10.0 ** (rnd(-10, 5) / 10.0)).contiguous()
        impl = rnd(0, 2)
        _lib.call("rl4co_set_rollout_impl", impl)
        try:
            res = {}
            for mode in ("never", "always"):
                pol.first_context_in_kernel = mode
                res[mode] = pol(td, env, decode_type="greedy", hidden=h, return_actions=True, return_sum_log_likelihood=False)
            a5, a4 = res["never"]["actions"], res["always"]["actions"]
            ok = True
            diff = (a5 != a4).any(1).nonzero().flatten()
            if diff.numel():
                # at the first differing step both paths saw the same state: the two chosen nodes must be a near tie
                ev5 = pol(td[diff], env, actions=a5[diff], hidden=h[diff], return_sum_log_likelihood=False)["log_likelihood"]
                ev4 = pol(td[diff], env, actions=a4[diff], hidden=h[diff], return_sum_log_likelihood=False)["log_likelihood"]
                first = (a5[diff] != a4[diff]).float().argmax(1, keepdim=True)
                gap = (ev5.gather(1, first) - ev4.gather(1, first)).abs().max()
                ok &= bool(gap < 1e-4)
            same = ~(a5 != a4).any(1)
            ll5, ll4 = res["never"]["log_likelihood"].sum(1), res["always"]["log_likelihood"].sum(1)
            ok &= bool(torch.allclose(ll4[same], ll5[same], rtol=1e-5, atol=2e-5))
            ok &= bool(torch.equal(a4.sort(1).values, torch.arange(num_loc, device=dev).expand(B, num_loc)))
            acts = torch.rand(B, num_loc, device=dev).argsort(1)
            for mode in ("never", "always"):
                pol.first_context_in_kernel = mode
                res[mode] = pol(td, env, actions=acts, hidden=h)["log_likelihood"]
            ok &= bool(torch.allclose(res["always"], res["never"], rtol=1e-5, atol=2e-5))
        finally:
            _lib.call("rl4co_set_rollout_impl", 0)
        if not ok:
            bad += 1
            print(f"FIRSTCTX FAIL tsp{num_loc} B={B} impl={impl}: {int(diff.numel())} tours differ, "
                  f"evaluate diff {float((res['always'] - res['never']).abs().max()):.2e}", flush=True)
        n += 1
print(f"first-node context: {n} random cases, {bad - bad_prev} failures", flush=True)
sys.exit(1 if bad else 0)
