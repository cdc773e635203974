"""C5 (SLAP, examples/slap.py wiring): the constant-query kernel (am_rollout_slap.cu, impl 0) against the generic
register-file kernel (impl 2) - CUDA-event timings and a cross-check of greedy / sampled / evaluated rollouts.
    python tools/time_slap.py [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib

dev = torch.device("cuda:0")
Bs = [int(x) for x in sys.argv[1:]] or [64, 16384, 131072]
EMBED = 128


def timed(fn, warm=2, reps=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


with torch.inference_mode():
    for (na, nl) in ((10, 10), (5, 10), (3, 7)):
        N = na * nl
        torch.manual_seed(1234)
        env = R.SLAPEnv(generator_params={"n_aisles": na, "n_locs": nl}, device=dev)
        pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
        for B in (Bs if N == 100 else Bs[:1]):
            td = env.reset(env.generator(B).to(dev))
            h = torch.cat([pol.encoder(td[i:i + 8192])[0] for i in range(0, B, 8192)])
            res = {}
            for impl, tag in ((0, "slap"), (2, "rf")):
                _lib.call("rl4co_set_rollout_impl", impl)
                for mode in ("greedy", "sampling"):
                    ms = timed(lambda: pol(td, env, decode_type=mode, hidden=h))
                    torch.manual_seed(7)
                    o = pol(td, env, decode_type=mode, hidden=h, return_actions=True)
                    res[(tag, mode)] = o
                    algo = 3 * N * EMBED * 4 + EMBED * 4 + N * 8 + 20 * 5 * 8 + (20 * 8 + 20 * 4 + 8)
                    print(f"slap{N} B={B} {tag:5s} {mode:9s} {ms:9.3f} ms  {B / ms * 1e-3:7.3f} M rollouts/s  "
                          f"{B * algo / ms * 1e-6:7.0f} GB/s  mean reward {o['reward'].mean().item():.4f}", flush=True)
            for mode in ("greedy", "sampling"):
                ga, gb = res[("slap", mode)], res[("rf", mode)]
                same = (ga["actions"] == gb["actions"]).all(1)
                if not bool(same.any()):
                    print(f"   {mode}: assignments equal 0/{same.numel()}")
                    continue
                print(f"   {mode}: assignments equal {int(same.sum())}/{same.numel()}  max |ll diff| on equal "
                      f"{(ga['log_likelihood'] - gb['log_likelihood'])[same].abs().max().item():.2e} (|ll| up to "
                      f"{gb['log_likelihood'].abs().max().item():.1f})  reward diff "
                      f"{(ga['reward'] - gb['reward'])[same].abs().max().item():.2e}")
            # evaluate: the rf kernel's sampled assignments scored by both kernels
            acts = res[("rf", "sampling")]["actions"]
            ev = {}
            for impl, tag in ((0, "slap"), (2, "rf")):
                _lib.call("rl4co_set_rollout_impl", impl)
                ev[tag] = pol(td, env, actions=acts, hidden=h)
            _lib.call("rl4co_set_rollout_impl", 0)
            d = (ev["slap"]["log_likelihood"] - ev["rf"]["log_likelihood"]).abs().max().item()
            d2 = (ev["rf"]["log_likelihood"] - res[("rf", "sampling")]["log_likelihood"]).abs().max().item()
            print(f"   evaluate: max |ll diff| slap vs rf {d:.2e}; rf evaluate vs rf sampling {d2:.2e}")
