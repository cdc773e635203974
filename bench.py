#!/usr/bin/env python
"""Benchmark of the rollout path: TSP100 greedy Attention-Model rollouts/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--batch B_per_gpu] [--no-cpu-baseline] [--no-e2e]

One "step" = one pass of the hot path over one batch of synthetic TSP100 instances
(BASELINE.json configs[2]: 2^20 instances per GPU, uniform coordinates, random-init AM
policy of the reference's default shape).

`value`     decode path (K3 projection + K2 persistent rollout kernel + reward-statistics
            reduction) with the encoder output already resident in HBM - the path the
            north star names; rollouts/s summed over all ranks (weak scaling).
`e2e`       the same metric through the public API `policy(td, env)` from PINNED HOST
            coordinates: H2D copy + PyTorch encoder + K3 + K2 + D2H of the rewards.
`roofline`  K2 kernel alone (CUDA events around its launches) vs the measured HBM peak.
`cpu_baseline` / `--impl reference`: the reference algorithm on the host cores (PyTorch-CPU
            encoder + oracle/port.py decode loop, all threads) on a bounded sample.
`side_configs` (N = 1 only) the other four BASELINE.json configurations, each with its own clock
            sample, roofline object and host-core number: C1 TSP50 B=512 greedy (both arms on that
            exact config), C2 TSP100 env-only replay B=2^20, C4 CVRP100 POMO multistart sampling,
            C5 SLAP greedy batch sweep.
`parity_check` 256 instances of the timed batch against oracle/port.py (outside the timed region).
`strong_scaling` (N > 1 only) BASELINE configs[2] read literally: ONE 2^20 batch split over the ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_NODES, EMBED = 100, 128
# algorithmic bytes of one TSP-N AM greedy rollout (fp32 caches read once,
# tour + reward + ll written): 16NE + 4E + 16N + 8
ALGO_BYTES_PER_ROLLOUT = 16 * N_NODES * EMBED + 4 * EMBED + 16 * N_NODES + 8   # 206 920 B


def profile_summary():
    """ncu-derived per-launch numbers of the dominant kernel (profiles/r01_summary.json, written
    from the committed ncu capture by tools/summarize_profile.py); None when absent."""
    for name in ("r02_summary.json", "r01_summary.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                out = json.load(f)
            out["_file"] = f"profiles/{name}"
            return out
        except Exception:  # noqa: BLE001
            continue
    return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_policy(device):
    import rl4co_slap_b200 as R
    torch.manual_seed(1234)  # the reference's default seed (configs/experiment/routing/am.yaml:41)
    pol = R.AttentionModelPolicy(env_name="tsp", embed_dim=128, num_heads=8, num_encoder_layers=3,
                                 normalization="batch", feedforward_hidden=512)
    pol = pol.to(device).eval()
    if os.environ.get("RL4CO_BENCH_FIVE_TABLES"):   # A/B hook: project table 4 instead of forming its one row in K2
        pol.first_context_in_kernel = "never"
    return pol


# ------------------------------------------------------------------------------------------
def cpu_reference(policy_cpu, batch: int, steps: int, warmup: int, threads: int):
    """Reference algorithm on the host: PyTorch-CPU encoder (rl4co/models/zoo/am/encoder.py
    math) + oracle/port.py greedy decode loop (the reference's op order).  rollouts/s."""
    from oracle import port
    torch.set_num_threads(threads)
    sd = policy_cpu.state_dict()
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"],
         "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"],
         "Wctx": sd["decoder.context_embedding.project_context.weight"],
         "Wph": sd["decoder.context_embedding.W_placeholder"]}
    g = torch.Generator().manual_seed(1234)
    times = []
    with torch.inference_mode():
        for it in range(warmup + steps):
            locs = torch.rand(batch, N_NODES, 2, generator=g)
            t0 = time.perf_counter()
            h, _ = policy_cpu.encoder({"locs": locs})
            out = port.am_rollout("tsp", port.tsp_reset(locs), h, W, "greedy")
            float(out["reward"].mean())
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    return batch / (sum(times) / len(times)), sum(times) / len(times)


def oracle_weights_of(policy):
    """Decoder weights of a product policy in the dict layout oracle/port.py takes (CPU tensors)."""
    sd = {k: v.detach().float().cpu() for k, v in policy.state_dict().items()}
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"], "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"]}
    for key, name in (("Wctx", "decoder.context_embedding.project_context.weight"),
                      ("Wph", "decoder.context_embedding.W_placeholder")):
        if name in sd:
            W[key] = sd[name]
    return W


def timed_region(fn, dev_index, min_seconds=1.0, warmup=3, max_reps=4000):
    """CUDA-event time of `fn` (ms per call) over a region of at least `min_seconds`, with the clock
    sampler running over exactly that region.  -> (ms_per_call, reps, clocks)"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    est = max(e0.elapsed_time(e1) * 1e-3, 1e-6)
    reps = int(min(max_reps, max(3, -(-min_seconds // est))))
    with ClockSampler(dev_index) as clk:
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if e0.elapsed_time(e1) < 350.0:      # let nvidia-smi deliver at least a few samples of this region
            time.sleep(0.35)
    return e0.elapsed_time(e1) / reps, reps, clk.summary()


def best_of(fn, reps=3, warmup=1):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return min(ts)


def side_configs(dev, peak_hbm, peak_src, quick=False):
    """BASELINE.json configs[0], [1], [3], [4] (the headline line is configs[2]).  Every entry: the GPU arm
    (CUDA events, own clock sample), its roofline against the bound SURVEY.md 8(d) names for it, and the
    reference algorithm (oracle/port.py + PyTorch-CPU encoder) on the host cores beside it."""
    import rl4co_slap_b200 as R
    from oracle import port
    from rl4co_slap_b200 import _lib

    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    idx = dev.index or 0
    fp32_peak = 148 * 128 * 2 * 1.965e9 / 1e12      # TFLOP/s, nominal CUDA-core FMA peak (not in MEASURED_PEAKS.json)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak_tensor = float(json.load(f)["bf16_tflops_sustained"])
        peak_tensor_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained; fp16 runs at the bf16 rate)"
    except Exception:  # noqa: BLE001
        peak_tensor, peak_tensor_src = 1590.0, "fallback (B200_PROFILING.md: 1.59 PFLOP/s dense bf16)"
    out = {}

    def pinned(x):
        return x.cpu().pin_memory()

    # ---- C1: TSPEnv num_loc=50, batch 512, greedy AM, both arms on exactly this config ---------------
    B, N = 512, 50
    pol = build_policy(dev)
    pol_cpu = build_policy("cpu")
    env = R.TSPEnv(generator_params=dict(num_loc=N), device=dev, check_solution=False)
    torch.manual_seed(1234)
    host_locs = torch.rand(B, N, 2).pin_memory()
    host_reward = torch.empty(B).pin_memory()
    dlocs = torch.empty(B, N, 2, device=dev)

    def c1_step():
        dlocs.copy_(host_locs, non_blocking=True)
        td = env.reset(R.TensorDict({"locs": dlocs}, batch_size=[B]))
        o = pol(td, env, decode_type="greedy")
        host_reward.copy_(o["reward"], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    ms, reps, clocks = timed_region(c1_step, idx, 0.6 if quick else 1.5)
    gpu_mean = float(host_reward.mean())
    # the same step with encoder + K3 + K2 captured as one CUDA graph (policy.graphed): B = 512 is launch bound
    with torch.inference_mode():
        td_static = env.reset(R.TensorDict({"locs": dlocs}, batch_size=[B]))
        fast = pol.graphed(td_static, env, decode_type="greedy")

    def c1_graph_step():
        dlocs.copy_(host_locs, non_blocking=True)       # dlocs IS the captured input buffer
        o = fast(td_static)
        host_reward.copy_(o["reward"], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    with torch.inference_mode():
        ms_graph, reps_graph, _ = timed_region(c1_graph_step, idx, 0.4 if quick else 1.0)
    graph_mean = float(host_reward.mean())
    W = oracle_weights_of(pol_cpu)
    ref = {}

    def c1_cpu():
        h, _ = pol_cpu.encoder({"locs": host_locs})
        ref["o"] = port.am_rollout("tsp", port.tsp_reset(host_locs), h, W, "greedy")

    with torch.inference_mode():
        sec = best_of(c1_cpu)
    algo = 16 * N * EMBED + 4 * EMBED + 16 * N + 8
    out["C1_tsp50_b512_greedy"] = {
        "workload": "TSPEnv num_loc=50, batch 512, greedy AttentionModel (3-layer encoder + decode), random-init weights "
                    "(BASELINE configs[0]); GPU arm end to end from pinned host coordinates to host rewards",
        "batch": B, "ms_per_step": ms, "reps": reps, "rollouts_per_s": B / (ms * 1e-3), "env_steps_per_s": B * N / (ms * 1e-3),
        "mean_reward": gpu_mean, "launches_per_step": pol.last_launches, "clocks": clocks,
        "cuda_graph": {"ms_per_step": ms_graph, "reps": reps_graph, "rollouts_per_s": B / (ms_graph * 1e-3),
                       "mean_reward": graph_mean,
                       "what": "policy.graphed(): the same end-to-end step (H2D, encoder, K3, K2, D2H) with the kernels "
                               "replayed as one CUDA graph instead of ~27 eager launches"},
        "roofline": {"bound": "hbm", "achieved": B * algo / (ms * 1e-3) / 1e9, "peak": peak_hbm, "unit": "GB/s",
                     "frac": B * algo / (ms * 1e-3) / 1e9 / peak_hbm, "traffic": None,
                     "note": "512 instances are 3.5 per SM: launch / latency bound, not a bandwidth case"},
        "cpu_baseline": {"value": B / sec, "unit": "rollouts/s", "cores": threads, "kind": "port",
                         "sample": f"the same 512 x TSP50 batch, best of 3 after 1 warm-up, {sec:.3f} s per pass",
                         "mean_reward": float(ref["o"]["reward"].mean()),
                         "tours_equal_to_gpu_arm": None}}
    with torch.inference_mode():
        td = env.reset(R.TensorDict({"locs": dlocs}, batch_size=[B]))
        o = pol(td, env, decode_type="greedy", return_actions=True)
        out["C1_tsp50_b512_greedy"]["cpu_baseline"]["tours_equal_to_gpu_arm"] = int(
            (o["actions"].cpu() == ref["o"]["actions"]).all(1).sum())
    del pol_cpu

    # ---- C2: TSP100 env-only replay of fixed random action sequences, batch 2^20 ---------------------
    B, N = (1 << 18) if quick else (1 << 20), 100
    env = R.TSPEnv(generator_params=dict(num_loc=N), device=dev, check_solution=True)
    torch.manual_seed(1234)
    locs = torch.rand(B, N, 2, device=dev)
    acts = torch.rand(B, N, device=dev).argsort(1)
    bufs = {"first": torch.empty(B, dtype=torch.int64, device=dev), "cur": torch.empty(B, dtype=torch.int64, device=dev),
            "i": torch.empty(B, 1, dtype=torch.int64, device=dev), "mask": torch.empty(B, N, dtype=torch.uint8, device=dev),
            "done": torch.empty(B, dtype=torch.uint8, device=dev), "reward": torch.empty(B, device=dev)}
    status = _lib.new_status(dev)
    stream = _lib.stream_ptr(dev)

    def c2_kernel():
        _lib.call("rl4co_tsp_replay", B, N, N, _lib.ptr(locs), _lib.ptr(acts), _lib.ptr(bufs["first"]), _lib.ptr(bufs["cur"]),
                  _lib.ptr(bufs["i"]), _lib.ptr(bufs["mask"]), _lib.ptr(bufs["done"]), _lib.ptr(bufs["reward"]), 1,
                  _lib.ptr(status), stream)

    ms, reps, clocks = timed_region(c2_kernel, idx, 0.5 if quick else 1.0)
    _lib.raise_for_status(status)
    td0 = R.TensorDict({"locs": locs}, batch_size=[B])
    ms_api, _, _ = timed_region(lambda: env.replay(td0, acts), idx, 0.3, max_reps=300)
    algo = 17 * N + 29
    nb = 65536 if not quick else 8192
    cl, ca = locs[:nb].cpu(), acts[:nb].cpu()

    def c2_cpu():
        st = port.tsp_reset(cl)
        for t in range(N):
            st["action"] = ca[:, t]
            st = port.tsp_step(st)
        port.tsp_reward(cl, ca, True)

    with torch.inference_mode():
        sec = best_of(c2_cpu, reps=2)
    out["C2_tsp100_env_replay"] = {
        "workload": "TSP100 env-only replay of fixed random permutations: reset + 100 x _step + get_reward + validity check "
                    "in one launch (BASELINE configs[1]); inputs 1.7 GB > L2",
        "batch": B, "ms_per_step": ms, "reps": reps, "rollouts_per_s": B / (ms * 1e-3), "env_steps_per_s": B * N / (ms * 1e-3),
        "api_ms_per_step": ms_api, "api_rollouts_per_s": B / (ms_api * 1e-3),
        "api_note": "env.replay(td, actions): the same launch + output allocation + one status-word D2H read",
        "mean_reward": float(bufs["reward"].mean()), "launches_per_step": 1, "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "tsp_replay (K1)", "achieved": B * algo / (ms * 1e-3) / 1e9, "peak": peak_hbm,
                     "unit": "GB/s", "frac": B * algo / (ms * 1e-3) / 1e9 / peak_hbm, "traffic": None,
                     "algorithmic_bytes_per_rollout": algo, "peak_source": peak_src},
        "cpu_baseline": {"value": nb / sec, "unit": "rollouts/s", "cores": threads, "kind": "port",
                         "sample": f"first {nb} instances of the same batch, step-by-step replay + reward + validity, "
                                   f"best of 2, {sec:.2f} s per pass"}}
    del locs, acts, bufs, td0
    torch.cuda.empty_cache()

    # ---- C4: CVRP100 POMO multistart sampling, S = 100 -------------------------------------------------
    N, S = 101, 100
    torch.manual_seed(1234)
    env = R.CVRPEnv(generator_params=dict(num_loc=100), device=dev, check_solution=False)
    pol = R.POMOPolicy(env_name="cvrp").to(dev).eval()
    c4 = {}
    for Bc in ((64,) if quick else (64, 4096)):
        with torch.inference_mode():
            td = env.reset(env.generator(Bc).to(dev))
            h = torch.cat([pol.encoder(td[i:i + 2048])[0] for i in range(0, Bc, 2048)])
            last = {}

            def c4_step():
                last["o"] = pol(td, env, decode_type="multistart_sampling", num_starts=S, hidden=h)

            ms, reps, clocks = timed_region(c4_step, idx, 0.5 if quick else 1.0, warmup=2, max_reps=200)
            ms_full, _, _ = timed_region(lambda: pol(td, env, decode_type="multistart_sampling", num_starts=S), idx, 0.3,
                                         warmup=1, max_reps=100)
            o = pol(td, env, decode_type="multistart_sampling", num_starts=S, hidden=h, return_actions=True)
            T = int(o["actions"].shape[1])
            mean_steps = float((o["actions"] != 0).sum(1).float().mean()) + 1
        macs = (129 * 128 + 3 * N * EMBED + EMBED * EMBED)      # per start and step, reference op order
        tf = Bc * S * T * macs * 2 / (ms * 1e-3) / 1e12
        c4[f"B{Bc}"] = {"batch": Bc, "starts": S, "steps": T, "mean_steps_until_done": mean_steps, "ms_per_step": ms, "reps": reps,
                        "rollouts_per_s": Bc * S / (ms * 1e-3), "env_steps_per_s": Bc * S * T / (ms * 1e-3),
                        "with_encoder_rollouts_per_s": Bc * S / (ms_full * 1e-3),
                        "mean_reward": float(o["reward"].mean()), "launches_per_step": pol.last_launches, "clocks": clocks,
                        "roofline": {"bound": "tensor", "kernel": "am_rollout_ms_tc_kernel<cvrp, sampling> (tcgen05 kind::f16, "
                                     "starts = TMEM lanes)", "achieved": tf, "peak": peak_tensor, "unit": "TFLOP/s",
                                     "frac": tf / peak_tensor, "traffic": None,
                                     "flops": "ALGORITHMIC: 2 x (129*128 + 3*N*E + E*E) MAC per start-step in the reference's op "
                                              "order, all S*B rollouts stepped until the slowest is done (T steps); the kernel "
                                              "issues 3 fp16 MMAs per product (hi/lo split) on 128 x 112 padded tiles",
                                     "peak_source": peak_tensor_src,
                                     "vs_nominal_fp32_fma_peak": tf / fp32_peak,
                                     "what_binds": "not the tensor pipe (ncu: sm__pipe_tensor_cycles_active 4.6 % with 64 of 148 SMs "
                                                   "busy, profiles/r02_c4_pomo_rollout_tc_raw.csv): per score element one MUFU ex2 "
                                                   "and one fp32->fp16 pack on the quarter-rate pipe, and tcgen05.ld at 64 B/clk/SM "
                                                   "(900 clk per 128 x 112 fp32 tile) - DESIGN.md section 3"}}
    pol_cpu = R.POMOPolicy(env_name="cvrp").eval()
    pol_cpu.load_state_dict({k: v.cpu() for k, v in pol.state_dict().items()})
    W = oracle_weights_of(pol_cpu)
    nb = 8 if not quick else 2
    g = env.generator(nb)
    gen = {k: g[k].cpu() for k in ("depot", "locs", "demand")}
    ref = {}

    def c4_cpu():
        st = port.cvrp_reset(gen["depot"], gen["locs"], gen["demand"], 1.0)
        h, _ = pol_cpu.encoder(st)
        ref["o"] = port.am_rollout("cvrp", st, h, W, "multistart_sampling", num_starts=S, use_graph_context=False,
                                   check_solution=False)

    with torch.inference_mode():
        sec = best_of(c4_cpu, reps=2)
    out["C4_cvrp100_pomo_multistart_sampling"] = {
        "workload": "CVRP100 POMO policy (6-layer encoder, instance norm, no graph context), multistart sampling, 100 starts "
                    "per instance (BASELINE configs[3]); timed with the encoder output resident, "
                    "with_encoder_rollouts_per_s = policy(td, env) including the encoder",
        **c4,
        "cpu_baseline": {"value": nb * S / sec, "unit": "rollouts/s", "cores": threads, "kind": "port",
                         "sample": f"{nb} instances x 100 starts incl. the encoder, best of 2, {sec:.2f} s per pass, "
                                   f"{int(ref['o']['actions'].shape[1])} steps"}}
    del td, h, o, last
    torch.cuda.empty_cache()

    # ---- C5: SLAP (examples/slap.py wiring) greedy, batch sweep -----------------------------------------
    N, P = 100, 20
    torch.manual_seed(1234)
    env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
    pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
    algo = 3 * N * EMBED * 4 + EMBED * 4 + N * 8 + 20 * 5 * 8 + (P * 8 + P * 4 + 8)
    sweep = {}
    for Bs in ((64, 16384) if quick else (64, 1024, 16384, 131072, 1 << 20)):
        with torch.inference_mode():
            td = env.reset(env.generator(Bs).to(dev))
            h = torch.empty(Bs, N, EMBED, device=dev)
            for lo in range(0, Bs, 8192):
                h[lo:lo + 8192] = pol.encoder(td[lo:lo + 8192])[0]
            ms, reps, clocks = timed_region(lambda: pol(td, env, decode_type="greedy", hidden=h), idx,
                                            0.4 if Bs < (1 << 20) else 1.0, warmup=2, max_reps=2000)
            # the call = K3 projection + graph context + K2 rollout per chunk of 16 384 instances; the roofline is the
            # dominant kernel's (am_rollout_slap_kernel), timed by the policy's own CUDA events around its launches
            pol.profile = []
            for _ in range(3):
                o = pol(td, env, decode_type="greedy", hidden=h)
            torch.cuda.synchronize()
            k_ms = sum(ev[1].elapsed_time(ev[2]) for _, ev in pol.profile) / 3
            p_ms = sum(ev[0].elapsed_time(ev[1]) for _, ev in pol.profile) / 3
            pol.profile = None
            gb = Bs * algo / (k_ms * 1e-3) / 1e9
            sweep[f"B{Bs}"] = {"batch": Bs, "ms_per_step": ms, "reps": reps, "rollouts_per_s": Bs / (ms * 1e-3),
                               "env_steps_per_s": Bs * P / (ms * 1e-3), "mean_reward": float(o["reward"].mean()),
                               "launches_per_step": pol.last_launches, "clocks": clocks,
                               "kernel_ms_per_step": k_ms, "projection_ms_per_step": p_ms,
                               "kernel_rollouts_per_s": Bs / (k_ms * 1e-3),
                               "roofline": {"bound": "hbm", "kernel": "am_rollout_slap_kernel<13> (K2 for SLAP, one launch per chunk)",
                                            "achieved": gb, "peak": peak_hbm, "unit": "GB/s",
                                            "frac": gb / peak_hbm, "traffic": None, "algorithmic_bytes_per_rollout": algo,
                                            "timed": "the kernel's launches alone (policy CUDA events), not the whole call"}}
            del td, h, o
            torch.cuda.empty_cache()
    pol_cpu = R.AttentionModelPolicy(env_name="slap").eval()
    pol_cpu.load_state_dict({k: v.cpu() for k, v in pol.state_dict().items()})
    W = oracle_weights_of(pol_cpu)
    nb = 512 if not quick else 64
    g = env.generator(nb)
    gen = {k: g[k].cpu() for k in ("locs", "freq", "assignment", "picklist", "depot_loc_dist")}

    def c5_cpu():
        st = port.slap_reset(gen["locs"], gen["assignment"].clone(), P, {k: gen[k] for k in ("freq", "picklist", "depot_loc_dist")})
        h, _ = pol_cpu.encoder(gen)
        port.am_rollout("slap", st, h, W, "greedy")

    with torch.inference_mode():
        sec = best_of(c5_cpu, reps=2)
    out["C5_slap_greedy_sweep"] = {
        "workload": "SLAPEnv 10 aisles x 10 locations, 20 products, 20 orders x 5 SKUs, greedy AM rollout with the "
                    "examples/slap.py embeddings (BASELINE configs[4]); encoder output resident, instances generated on the device",
        **sweep,
        "cpu_baseline": {"value": nb / sec, "unit": "rollouts/s", "cores": threads, "kind": "port",
                         "sample": f"{nb} instances incl. the encoder, best of 2, {sec:.2f} s per pass"}}
    return out


def parity_check(pol, env, locs, h, n=256):
    """Out of the timed region: `n` instances of the timed batch against oracle/port.py with the bench policy -
    greedy tours, and evaluate-mode ll / reward on the oracle's own tours (no argmax divergence)."""
    import rl4co_slap_b200 as R
    from oracle import port
    W = oracle_weights_of(pol)
    sub, hs = locs[:n].contiguous(), h[:n].contiguous()
    with torch.inference_mode():
        ref = port.am_rollout("tsp", port.tsp_reset(sub.cpu()), hs.cpu(), W, "greedy")
        td = env.reset(R.TensorDict({"locs": sub}, batch_size=[n]))
        ours = pol(td, env, decode_type="greedy", hidden=hs, return_actions=True)
        ev = pol(td, env, actions=ref["actions"].to(sub.device), hidden=hs)
    rel = lambda a, b: float(((a.cpu() - b).abs() / b.abs().clamp_min(1e-12)).max())
    return {"n": n, "tours_equal": int((ours["actions"].cpu() == ref["actions"]).all(1).sum()),
            "max_rel_ll": rel(ev["log_likelihood"], ref["log_likelihood"]),
            "max_rel_reward": rel(ev["reward"], ref["reward"]),
            "what": "greedy tours vs oracle/port.py on the same encoder output; ll and reward in evaluate mode on the oracle's tours"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    import rl4co_slap_b200 as R  # noqa: F401  (policy weights / encoder module only; CPU)
    pol = build_policy("cpu")
    batch = args.ref_batch
    v, sec = cpu_reference(pol, batch, max(1, args.steps), min(args.warmup, 1), threads)
    line = {"impl": "reference", "metric": "tsp100_greedy_am_rollouts_per_s", "value": v, "unit": "rollouts/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "env_steps_per_s": v * N_NODES,
            "config": {"workload": "TSP100 AttentionModel greedy decode (BASELINE configs[2])",
                       "sample": f"batch {batch} per step on the host"},
            "cpu_baseline": {"value": v, "unit": "rollouts/s", "cores": threads, "kind": "port",
                             "sample": f"TSP100 greedy AM (3-layer encoder + decode), batch {batch} x {args.steps} steps"},
            "e2e": {"value": v, "unit": "rollouts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist

    import rl4co_slap_b200 as R
    from rl4co_slap_b200 import baselines

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rollout path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    R._lib.load()

    B = args.batch
    pol = build_policy(dev)
    env = R.TSPEnv(generator_params=dict(num_loc=N_NODES), device=dev, check_solution=False)
    torch.manual_seed(1234 + rank)
    locs = torch.rand(B, N_NODES, 2, device=dev)
    td = env.reset(R.TensorDict({"locs": locs}, batch_size=[B]))

    # encoder output for the whole batch, resident in HBM (untimed setup; 51 GB at B = 2^20)
    h = torch.empty(B, N_NODES, EMBED, dtype=torch.float32, device=dev)
    with torch.inference_mode():
        for lo in range(0, B, 8192):
            h[lo:lo + 8192] = pol.encoder({"locs": locs[lo:lo + 8192]})[0]

    stats = torch.zeros(3, dtype=torch.float64, device=dev)

    def decode_step():
        out = pol(td, env, decode_type="greedy", hidden=h)
        stats.zero_()
        baselines.reward_stats(out["reward"], stats)         # K4: local partials ...
        baselines.all_reduce_sum_(stats)                      # ... one tiny NCCL all-reduce
        return out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, region=False):
        for _ in range(warmup):
            fn()
        barrier()
        if region:                                            # ncu --profile-from-start off: timed steps only
            torch.cuda.profiler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        if region:
            torch.cuda.profiler.stop()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() / steps

    with torch.inference_mode():
        decode_step()                                         # first touch (allocator, lib load)
        pol.profile = []
        with ClockSampler(local) as clk:
            ms_step = timed(decode_step, args.steps, args.warmup, region=args.profile_region)
        prof = pol.profile[-args.steps * ((B + pol.rollout_chunk - 1) // pol.rollout_chunk):]
        pol.profile = None
        launches_per_step = pol.last_launches + 1             # + reward_stats
        torch.cuda.synchronize()
        k2_ms = sum(ev[1].elapsed_time(ev[2]) for _, ev in prof)
        k3_ms = sum(ev[0].elapsed_time(ev[1]) for _, ev in prof)
        k2_rollouts = sum(n for n, _ in prof)
        mean_reward = (stats[0] / stats[2]).item()

        # ---- end to end through the public API from pinned host memory -------------------
        e2e = None
        if not args.no_e2e:
            host_locs = locs.cpu().pin_memory()
            host_reward = torch.empty(B, dtype=torch.float32).pin_memory()
            dlocs = torch.empty_like(locs)

            def e2e_step():
                dlocs.copy_(host_locs, non_blocking=True)
                tdx = env.reset(R.TensorDict({"locs": dlocs}, batch_size=[B]))
                out = pol(tdx, env, decode_type="greedy")
                host_reward.copy_(out["reward"], non_blocking=True)
                torch.cuda.current_stream().synchronize()

            ms_e2e = timed(e2e_step, args.steps, args.warmup, region=args.profile_e2e)
            e2e = {"value": B * world / (ms_e2e * 1e-3), "unit": "rollouts/s",
                   "h2d_bytes_per_step": host_locs.numel() * 4, "d2h_bytes_per_step": host_reward.numel() * 4,
                   "ms_per_step": ms_e2e,
                   "includes": "H2D locs + fp32 encoder (PyTorch modules whose linears, skip + batch-norm epilogues and "
                               "self-attention run on this repo's own tcgen05 kernels rl4co_linear_tc[_ex] / "
                               "rl4co_encoder_attention_tc - no cuBLAS / cuDNN) + K3 + K2 + D2H rewards"}

        # ---- strong scaling: ONE 2^20 batch split over the ranks (BASELINE configs[2] read literally) ----
        strong = None
        if world > 1 and not args.no_strong:
            Bs = max(1, (1 << 20) // world)
            td_s = env.reset(R.TensorDict({"locs": locs[:Bs]}, batch_size=[Bs]))
            h_s = h[:Bs]

            def strong_step():
                out = pol(td_s, env, decode_type="greedy", hidden=h_s)
                stats.zero_()
                baselines.reward_stats(out["reward"], stats)
                baselines.all_reduce_sum_(stats)

            ms_s = timed(strong_step, args.steps, args.warmup)
            strong = {"global_batch": Bs * world, "per_gpu_batch": Bs, "ms_per_step": ms_s,
                      "value": Bs * world / (ms_s * 1e-3), "unit": "rollouts/s", "scaling": "strong",
                      "timed_region": "as `value`: K3 + K2 + reward-stat all-reduce, encoder output resident"}

        # ---- parity of the benchmark configuration itself (outside every timed region) ------------------
        parity = parity_check(pol, env, locs, h, 256) if rank == 0 and not args.no_parity else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    prof_sum = profile_summary()
    traffic = None
    if prof_sum and prof_sum.get("k2_dram_bytes_per_rollout"):
        traffic = prof_sum["k2_dram_bytes_per_rollout"] * min(B, pol.rollout_chunk)   # per launch, like `achieved`
    achieved = k2_rollouts * ALGO_BYTES_PER_ROLLOUT / (k2_ms * 1e-3) / 1e9 if k2_ms > 0 else 0.0
    value = B * world / (ms_step * 1e-3)
    line = {
        "metric": "tsp100_greedy_am_rollouts_per_s", "value": value, "unit": "rollouts/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "env_steps_per_s": value * N_NODES,
        "config": {"workload": "TSP100 AttentionModel greedy decode (BASELINE configs[2])",
                   "per_gpu_batch": B, "global_batch": B * world, "num_loc": N_NODES, "embed_dim": EMBED,
                   "parallelism": f"dp{world}", "timed_region": "K3 projection + K2 rollout + reward-stat all-reduce; "
                   "encoder output resident in HBM", "l2": "inputs larger than L2 (encoder output "
                   f"{B * N_NODES * EMBED * 4 / 1e9:.1f} GB per GPU)", "rollout_chunk": pol.rollout_chunk,
                   "mean_reward": mean_reward},
        "roofline": {"bound": "hbm", "kernel": "am_rollout_tsp_greedy_kernel<7> (K2, one launch per chunk)",
                     "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_rollout": ALGO_BYTES_PER_ROLLOUT,
                     "algorithmic_bytes_per_launch": ALGO_BYTES_PER_ROLLOUT * min(B, pol.rollout_chunk),
                     "kernel_ms_per_step": k2_ms / args.steps, "projection_ms_per_step": k3_ms / args.steps,
                     "kernel_share_of_step": (k2_ms / args.steps) / ms_step,
                     # second ceiling: the folded decode step is 3*N*E fp32 MACs per step if no visited node
                     # were skipped (the compacted stages issue fewer); the denominators are the nominal 128 MAC/clk/SM
                     # and the 85 MAC/clk/SM this repo's own FFMA2 probe sustained (tools/probe_sm.cu) - a builder-side
                     # probe, not a driver-measured peak
                     "fp32_ceiling": {
                         "nominal_macs_per_rollout": N_NODES * 3 * N_NODES * EMBED,
                         "frac_of_nominal_128_mac_clk_sm": (k2_rollouts * N_NODES * 3 * N_NODES * EMBED) /
                                                           (k2_ms * 1e-3) / (128.0 * 148 * 1.965e9) if k2_ms > 0 else None,
                         "frac_of_probed_85_mac_clk_sm_ffma2": (k2_rollouts * N_NODES * 3 * N_NODES * EMBED) /
                                                               (k2_ms * 1e-3) / (85.0 * 148 * 1.965e9) if k2_ms > 0 else None,
                         "note": "counts the MACs of the un-compacted step; node compaction issues fewer"}},
        "clocks": clk.summary(),
        "gpu_launches": launches_per_step * args.steps,
    }
    line["roofline"]["traffic_source"] = (f"{prof_sum['_file']} (ncu --set full capture of this kernel, dram__bytes_read + "
                                          "dram__bytes_write per rollout x rollouts per launch; NOT measured in this run)"
                                          if traffic is not None else None)
    if e2e is not None:
        line["e2e"] = e2e
    if strong is not None:
        line["strong_scaling"] = strong
    if parity is not None:
        line["parity_check"] = parity
    if world == 1 and not args.no_side:
        del h, td, locs, pol, env
        if e2e is not None:
            del host_locs, host_reward, dlocs
        torch.cuda.empty_cache()
        line["side_configs"] = side_configs(dev, peak, peak_src, quick=args.quick_side)
    if not args.no_cpu_baseline and world == 1:     # host-core baseline: rank 0 of the single-GPU run only
        threads = os.cpu_count() or 1
        v, sec = cpu_reference(build_policy("cpu"), args.ref_batch, args.ref_steps, 1, threads)
        line["cpu_baseline"] = {"value": v, "unit": "rollouts/s", "cores": threads, "kind": "port",
                                "sample": f"TSP100 greedy AM (3-layer encoder + decode), {args.ref_steps} timed passes of "
                                          f"batch {args.ref_batch} after 1 warm-up, {sec:.2f} s per pass"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--batch", type=int, default=1 << 20, help="instances per GPU")
    ap.add_argument("--ref-batch", type=int, default=4096, help="host-side sample batch")
    ap.add_argument("--ref-steps", type=int, default=4, help="timed host-side passes of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-side", action="store_true", help="skip the side_configs block (C1, C2, C4, C5)")
    ap.add_argument("--quick-side", action="store_true", help="side_configs at reduced batch sizes (debugging)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling measurement at N > 1")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity_check block")
    ap.add_argument("--profile-e2e", action="store_true",
                    help="cudaProfilerStart/Stop around the timed end-to-end steps instead")
    ap.add_argument("--profile-region", action="store_true",
                    help="cudaProfilerStart/Stop around the timed decode steps (ncu --profile-from-start off)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
