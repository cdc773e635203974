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
    try:
        with open(os.path.join(ROOT, "profiles", "r01_summary.json")) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
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
    return pol.to(device).eval()


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
                   "ms_per_step": ms_e2e, "includes": "H2D locs + fp32 encoder (PyTorch modules; linears, skip+BN and attention on library kernels) + K3 + K2 + D2H rewards"}

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
                     # second ceiling: the folded decode step is 3*N*E fp32 MACs; FFMA2 sustains
                     # ~85 MAC/clk/SM on this part (tools/probe_sm.cu), the kernel is latency bound below that
                     "fma_frac_of_measured_ffma2_peak": (k2_rollouts * N_NODES * 3 * N_NODES * EMBED) /
                                                        (k2_ms * 1e-3) / (85.0 * 148 * 1.965e9) if k2_ms > 0 else None},
        "clocks": clk.summary(),
        "gpu_launches": launches_per_step * args.steps,
    }
    if e2e is not None:
        line["e2e"] = e2e
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
