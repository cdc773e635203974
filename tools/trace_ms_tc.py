"""Cycle trace of the tensor-core multistart kernel (csrc/am_rollout_ms_tc.cu built with -DMS_TRACE).
    python tools/trace_ms_tc.py build      # here (nvcc, no GPU): tools/_trace/ms/librl4co_trace.so
    python tools/trace_ms_tc.py [tsp|cvrp] # on the GPU box"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TRACE_LIB = os.path.join(ROOT, "tools", "_trace", "ms", "librl4co_trace.so")
if len(sys.argv) > 1 and sys.argv[1] == "build":
    from rl4co_slap_b200 import build
    print(build.build(extra_flags=["-DMS_TRACE"], lib=TRACE_LIB, obj_dir=os.path.dirname(TRACE_LIB)))
    sys.exit(0)
import numpy as np, torch
from rl4co_slap_b200 import _lib
_lib.LIB_PATH = TRACE_LIB
import rl4co_slap_b200 as R

name = sys.argv[1] if len(sys.argv) > 1 else "cvrp"
dev = torch.device("cuda:0")
torch.manual_seed(1234)
B, S = 148, 100
env = R.get_env(name, generator_params=dict(num_loc=100), device=dev, check_solution=False)
pol = R.POMOPolicy(env_name=name).to(dev).eval()
with torch.inference_mode():
    td = env.reset(env.generator(B).to(dev))
    h = pol.encoder(td)[0].float().contiguous()
N = h.shape[1]
T = N if name == "tsp" else 2 * (N - 1) + 1
W, q0, wcap, Wfixed = pol.decoder.folded_weights()
nt = W.shape[0] // 128
stream = _lib.stream_ptr(dev)
tables = torch.empty(nt, B * N, 128, device=dev)
_lib.call("rl4co_am_project_tc", B * N, nt, _lib.ptr(h), _lib.ptr(pol.decoder.weight_image()), _lib.ptr(tables), stream)
trace = torch.zeros(T * 16, dtype=torch.int64, device=dev)
R_ = B * S
acts = torch.empty(R_, T, dtype=torch.int64, device=dev)
logp = torch.empty(R_, T, device=dev)
ll, rew = torch.empty(R_, device=dev), torch.empty(R_, device=dev)
steps = torch.empty(R_, dtype=torch.int32, device=dev)
status = _lib.new_status(dev)
a = _lib.AmRolloutArgs()
a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = _lib.ENV_IDS[name], _lib.MODE_SAMPLING, B, S, 1, 100
a.N, a.T, a.P, a.O, a.K, a.n_tables = N, T, 0, 0, 0, nt
a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
a.tables, a.q0, a.wcap, a.locs = _lib.ptr(tables), _lib.ptr(q0), _lib.ptr(wcap), _lib.ptr(td["locs"].contiguous())
if name == "cvrp":
    a.demand, a.vehicle_capacity = _lib.ptr(td["demand"].contiguous()), _lib.ptr(td["vehicle_capacity"].contiguous())
a.seed, a.offset = 7, 0
a.actions, a.logp, a.logp_all, a.ll, a.reward = _lib.ptr(acts), _lib.ptr(logp), _lib.ptr(trace), _lib.ptr(ll), _lib.ptr(rew)
a.steps, a.status = _lib.ptr(steps), _lib.ptr(status)
for _ in range(2):
    trace.zero_()
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
torch.cuda.synchronize()
tr = trace.cpu().numpy().reshape(T, 16)
n = int(steps[0].item())
tr = tr[1:n]                       # steps 1 .. n-1 of rollout 0 (step 0 is the forced start)
d = np.diff(tr[:, :8], axis=1)
names = ["query", "wait scores(0)", "8 heads", "last glimpse", "wait logits", "logit phase", "record+transition"]
print(f"{name}100 S={S}: rollout 0 ran {n} steps, status {int(status.item())}, mean reward {rew.mean().item():.4f}")
m = d.mean(0)
print("per step: " + "  ".join(f"{k} {v:.0f}" for k, v in zip(names, m)))
print(f"   inside the heads: waiting for scores {tr[:, 8].mean():.0f}  waiting for P V {tr[:, 9].mean():.0f}  softmax arithmetic {tr[:, 10].mean():.0f}")
if tr[:, 11].any():
    q = tr[:, [0, 11, 12, 13, 14, 1]]
    print("   inside the query phase: " + "  ".join(f"{k} {v:.0f}" for k, v in zip(
        ["shuffles + loads issued", "loads landed + staged", "two passes of math + tcgen05.st", "wait::st", "fence + arrive"], np.diff(q, axis=1).mean(0))))
print("step start to next step start:", (tr[1:, 0] - tr[:-1, 0]).mean())
