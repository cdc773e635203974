"""TSP100 / CVRP100 POMO multistart timing (policy call with a precomputed encoder output): python tools/time_pomo.py [tsp|cvrp] [B] [S]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib

name = sys.argv[1] if len(sys.argv) > 1 else "tsp"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
S = int(sys.argv[3]) if len(sys.argv) > 3 else 100
dev = torch.device("cuda:0")
torch.manual_seed(0)
env = R.get_env(name, generator_params=dict(num_loc=100), device=dev)
pol = R.POMOPolicy(env_name=name).to(dev).eval()
td = env.reset(batch_size=[B])
with torch.inference_mode():
    h = pol.encoder(td)[0]
    for impl in (0, 2):
        _lib.call("rl4co_set_rollout_impl", impl)
        for dt in ("multistart_greedy", "multistart_sampling"):
            f = lambda: pol(td, env, decode_type=dt, num_starts=S, hidden=h)
            for _ in range(2):
                f()
            ts = []
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); o = f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            ms = sorted(ts)[1]
            print(f"{name}100 POMO B={B} S={S} impl={impl} {dt}: {ms:.2f} ms  {B * S / ms * 1e3 / 1e6:.3f} M rollouts/s  "
                  f"mean reward {o['reward'].mean().item():.4f}")
_lib.call("rl4co_set_rollout_impl", 0)
