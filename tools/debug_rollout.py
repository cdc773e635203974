"""Step-by-step smoke of the rollout kernels on the fixtures (prints before every launch so a
hang is attributable).  usage: python -u tools/debug_rollout.py case [impl]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib
from tests.helpers import env_of, fixture, gen_td, load_policy, t

case = sys.argv[1]
impl = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda:0")
_lib.call("rl4co_set_rollout_impl", impl)
fx, name = fixture(case), env_of(case)
pol, env = load_policy(name, dev), R.get_env(name, device=dev)
td = env.reset(gen_td(name, fx, dev))
for mode in ("greedy", "evaluate", "sampling"):
    print("launch", case, impl, mode, flush=True)
    kw = dict(actions=t(fx["greedy.actions"], dev)) if mode == "evaluate" else dict(decode_type=mode)
    out = pol(td, env, return_actions=True, hidden=t(fx["hidden"], dev), **kw)
    torch.cuda.synchronize()
    print("  done; reward", out["reward"][:4].tolist(), flush=True)
    if mode != "sampling":
        same = np.array_equal(out["actions"].cpu().numpy(), fx["greedy.actions"])
        print("  tours equal reference:", same, "| max ll diff", float(np.abs(out["log_likelihood"].cpu().numpy() - fx["greedy.ll"]).max()), flush=True)
