"""K2-only timing (CUDA events) of the single-start rollout kernels on random TSP instances.

    python tools/time_k2.py [N] [instances per SM] [impl] [greedy|sampling]     # impl: 0 auto, 2 generic kernel
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from rl4co_slap_b200 import _lib
import rl4co_slap_b200 as R
if os.environ.get("RL4CO_LIB"):          # A/B timing of a side build (tools/_exp/)
    _lib.LIB_PATH = os.environ["RL4CO_LIB"]

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
per_sm = int(sys.argv[2]) if len(sys.argv) > 2 else 64
impl = int(sys.argv[3]) if len(sys.argv) > 3 else 0
mode = {"greedy": _lib.MODE_GREEDY, "sampling": _lib.MODE_SAMPLING}[sys.argv[4] if len(sys.argv) > 4 else "greedy"]
dev = torch.device("cuda:0")
torch.manual_seed(1234)
B, T = 148 * per_sm, N
pol = R.AttentionModelPolicy(env_name="tsp").to(dev).eval()
locs = torch.rand(B, N, 2, device=dev)
with torch.inference_mode():
    h = pol.encoder({"locs": locs})[0].float().contiguous()
W, q0, wcap, Wfixed = pol.decoder.folded_weights()
nt = W.shape[0] // 128
stream = _lib.stream_ptr(dev)
tables = torch.empty(nt, B * N, 128, device=dev)
gctx = torch.empty(B, 128, device=dev)
_lib.call("rl4co_am_project", B * N, nt, _lib.ptr(h), _lib.ptr(W), _lib.ptr(tables), stream)
_lib.call("rl4co_am_graph_context", B, N, _lib.ptr(h), _lib.ptr(Wfixed), _lib.ptr(gctx), stream)
acts = torch.empty(B, T, dtype=torch.int64, device=dev)
logp = torch.empty(B, T, device=dev)
ll, rew = torch.empty(B, device=dev), torch.empty(B, device=dev)
steps = torch.empty(B, dtype=torch.int32, device=dev)
status = _lib.new_status(dev)
a = _lib.AmRolloutArgs()
a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = 0, mode, B, 0, 0, N
a.seed, a.offset = 1234, 0
a.N, a.T, a.P, a.O, a.K, a.n_tables = N, T, 0, 0, 0, nt
a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
a.tables, a.gctx, a.q0, a.locs = _lib.ptr(tables), _lib.ptr(gctx), _lib.ptr(q0), _lib.ptr(locs)
a.actions, a.logp, a.logp_all, a.ll, a.reward = _lib.ptr(acts), _lib.ptr(logp), None, _lib.ptr(ll), _lib.ptr(rew)
a.steps, a.status = _lib.ptr(steps), _lib.ptr(status)
_lib.call("rl4co_set_rollout_impl", impl)
ts = []
for i in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts[1:])[len(ts[1:]) // 2]
clk = ms * 1e-3 * 1.965e9 / per_sm
print(f"N={N} impl={impl} mode={mode} B={B}: {ms:.3f} ms  {B / ms * 1e3 / 1e6:.3f} M rollouts/s  {clk:.0f} clk/instance  {clk / T:.0f} clk/step"
      f"  mean reward {rew.mean().item():.5f} ll {ll.mean().item():.5f} status {int(status.item())}")
