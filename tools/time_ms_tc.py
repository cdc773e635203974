"""C4 (CVRP100 POMO, S = 100) and TSP100 POMO: tensor-core multistart kernel (impl 0) against the register-file SIMT
kernel (impl 2) - CUDA-event timings and a cross-check of the greedy tours / sampled log-likelihoods.
    python tools/time_ms_tc.py [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib

dev = torch.device("cuda:0")
Bs = [int(x) for x in sys.argv[1:]] or [64, 1024]


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
    for name in ("cvrp", "tsp"):
        torch.manual_seed(1234)
        env = R.get_env(name, generator_params=dict(num_loc=100), device=dev, check_solution=False)
        pol = R.POMOPolicy(env_name=name).to(dev).eval()
        for B in Bs:
            td = env.reset(env.generator(B).to(dev))
            h = torch.cat([pol.encoder(td[i:i + 2048])[0] for i in range(0, B, 2048)])
            res = {}
            for impl, tag in ((0, "tc"), (2, "simt")):
                _lib.call("rl4co_set_rollout_impl", impl)
                for mode in ("multistart_greedy", "multistart_sampling"):
                    ms = timed(lambda: pol(td, env, decode_type=mode, num_starts=100, hidden=h))
                    o = pol(td, env, decode_type=mode, num_starts=100, hidden=h, return_actions=True)
                    res[(tag, mode)] = (ms, o)
                    T = o["actions"].shape[1]
                    print(f"{name}100 B={B} {tag:5s} {mode:20s} {ms:9.3f} ms  {B * 100 / ms * 1e-3:7.3f} M rollouts/s  "
                          f"T={T}  mean reward {o['reward'].mean().item():.4f}", flush=True)
            _lib.call("rl4co_set_rollout_impl", 0)
            ga, gb = res[("tc", "multistart_greedy")][1], res[("simt", "multistart_greedy")][1]
            Tm = max(ga["actions"].shape[1], gb["actions"].shape[1])
            pa = torch.nn.functional.pad(ga["actions"], (0, Tm - ga["actions"].shape[1]))
            pb = torch.nn.functional.pad(gb["actions"], (0, Tm - gb["actions"].shape[1]))
            same = (pa == pb).all(1)
            print(f"   greedy tours equal: {int(same.sum())}/{same.numel()}  max |ll diff| on equal tours "
                  f"{(ga['log_likelihood'] - gb['log_likelihood'])[same].abs().max().item():.2e}  "
                  f"max rel reward diff {((ga['reward'] - gb['reward'])[same].abs() / gb['reward'][same].abs()).max().item():.2e}")
            if B <= 64:     # the sampled tours of the tensor-core kernel re-scored by the SIMT kernel (evaluate mode, single start rows)
                sa = res[("tc", "multistart_sampling")][1]
                _lib.call("rl4co_set_rollout_impl", 2)
                ev = pol(R.batchify(td, 100), env, actions=sa["actions"], hidden=R.batchify(h, 100))
                _lib.call("rl4co_set_rollout_impl", 0)
                print(f"   sampled tours re-evaluated by the SIMT kernel: max |ll diff| "
                      f"{(ev['log_likelihood'] - sa['log_likelihood']).abs().max().item():.2e} (|ll| up to "
                      f"{sa['log_likelihood'].abs().max().item():.0f}), reward {(ev['reward'] - sa['reward']).abs().max().item():.2e}")
