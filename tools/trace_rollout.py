"""Cycle trace of the register-file rollout kernel (csrc/am_rollout_rf.cu built with -DRF_TRACE).

    python tools/trace_rollout.py build      # here (nvcc, no GPU): tools/_trace/librl4co_trace.so
    python tools/trace_rollout.py [env]      # on the GPU box: per-phase cycles of one trajectory

Thread 0 of CTA 0 stamps clock64() at eight points of every step of its first instance:
0 step start | 1 scores + head max | 2 glimpse reduced (after __syncwarp) | 3 partial logits stored |
4 past barrier 1 | 5 slice done | 6 past barrier 2 | 7 selection + transition done.
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TRACE_LIB = os.path.join(ROOT, "tools", "_trace", "librl4co_trace.so")

ABL = os.environ.get("RF_ABL", "0")          # timing-only ablation bits (csrc/am_rollout_rf.cu)
if ABL != "0":
    TRACE_LIB = os.path.join(ROOT, "tools", "_trace", f"abl{ABL}", "librl4co_trace.so")
if os.environ.get("RF_VARIANT"):            # a side build with RF_EXTRA_FLAGS, e.g. RF_VARIANT=nopf RF_EXTRA_FLAGS=-DRL4CO_NO_H_PREFETCH
    TRACE_LIB = os.path.join(ROOT, "tools", "_trace", os.environ["RF_VARIANT"], "librl4co_trace.so")
if len(sys.argv) > 1 and sys.argv[1] == "build":
    from rl4co_slap_b200 import build
    extra = os.environ.get("RF_EXTRA_FLAGS", "").split()
    print(build.build(extra_flags=["-DRF_TRACE", f"-DRF_ABL={ABL}"] + extra, lib=TRACE_LIB, obj_dir=os.path.dirname(TRACE_LIB)))
    sys.exit(0)

import numpy as np, torch
from rl4co_slap_b200 import _lib
_lib.LIB_PATH = TRACE_LIB
import rl4co_slap_b200 as R

dev = torch.device("cuda:0")
torch.manual_seed(1234)
N = int(os.environ.get("RF_N", "100"))
B, T = 148 * 2, N
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
trace = torch.zeros((T + 3) * 8, dtype=torch.int64, device=dev)
acts = torch.empty(B, T, dtype=torch.int64, device=dev)
logp = torch.empty(B, T, device=dev)
ll, rew = torch.empty(B, device=dev), torch.empty(B, device=dev)
steps = torch.empty(B, dtype=torch.int32, device=dev)
status = _lib.new_status(dev)
a = _lib.AmRolloutArgs()
a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = 0, _lib.MODE_GREEDY, B, 0, 0, N
a.N, a.T, a.P, a.O, a.K, a.n_tables = N, T, 0, 0, 0, nt
a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
a.tables, a.gctx, a.q0, a.locs = _lib.ptr(tables), _lib.ptr(gctx), _lib.ptr(q0), _lib.ptr(locs)
a.actions, a.logp, a.logp_all, a.ll, a.reward = _lib.ptr(acts), _lib.ptr(logp), _lib.ptr(trace), _lib.ptr(ll), _lib.ptr(rew)
a.steps, a.status = _lib.ptr(steps), _lib.ptr(status)
if os.environ.get("RF_FC"):     # table 4 left out: the kernel forms the first node's context row (ABI 4: h / w_first)
    w_first = W[512:640]
    a.n_tables, a.h, a.w_first = 4, _lib.ptr(h), _lib.ptr(w_first)
    tables[4].fill_(float("nan"))
for _ in range(2):
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
torch.cuda.synchronize()
full = trace.cpu().numpy().reshape(T + 3, 8)
tr = full[:T]
d = np.diff(tr, axis=1)
names = ["q+scores+max", "softmax+V+butterfly", "L+store", "barrier1", "slice", "barrier2", "select+transition"]
print(f"N={N} ablation bits={ABL} mean reward", rew.mean().item())
for lo, hi in ((1, T // 2), (T // 2, T - 1)):
    m = d[lo:hi].mean(0)
    print(f"steps {lo}-{hi}: " + "  ".join(f"{n} {v:.0f}" for n, v in zip(names, m)), "| step total", (tr[lo + 1:hi + 1, 0] - tr[lo:hi, 0]).mean())
print("cycles from step start to next step start:", " ".join(str(int(v)) for v in (tr[1:, 0] - tr[:-1, 0])))
for r in (T, T + 1):
    if full[r, 0]:
        print(f"compaction {r - T}: ranks+node table {full[r,1]-full[r,0]}  K {full[r,2]-full[r,1]}  V {full[r,3]-full[r,2]}  L {full[r,4]-full[r,3]}")
if full[T + 2, 0]:
    v = full[T + 2]
    print(f"instance: table loads issued {v[1]-v[0]}  smem init {v[2]-v[1]}  context wait {v[3]-v[2]}  decode {v[4]-v[3]}  log-probs {v[5]-v[4]}  reward+outputs {v[6]-v[5]}  total {v[6]-v[0]}")
print("whole trajectory cycles", tr[-1, 7] - tr[0, 0])
