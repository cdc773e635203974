"""Cycle trace of the SLAP rollout kernel (csrc/am_rollout_slap.cu built with -DSLAPK_TRACE).

    python tools/trace_slap.py build      # here (nvcc, no GPU): tools/_trace_slap/librl4co_trace.so
    python tools/trace_slap.py            # on the GPU box: per-phase cycles of one trajectory

Thread 0 of CTA 0 stamps clock64() during its second instance: per step 0 start | 1 logits of the own nodes (butterfly
done) | 2 warp winner stored | 3 past the barrier | 4 block winner + transition + glimpse update done; per instance
0 start | 1 scores stored | 2 small state staged | 3 value table landed + first glimpse state | 4 decode done |
5 log-probs | 6 reward + outputs.
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TRACE_LIB = os.path.join(ROOT, "tools", "_trace_slap", "librl4co_trace.so")
if len(sys.argv) > 1 and sys.argv[1] == "build":
    from rl4co_slap_b200 import build
    print(build.build(extra_flags=["-DSLAPK_TRACE"], lib=TRACE_LIB, obj_dir=os.path.dirname(TRACE_LIB)))
    sys.exit(0)

import numpy as np, torch
from rl4co_slap_b200 import _lib
_lib.LIB_PATH = TRACE_LIB
import rl4co_slap_b200 as R

dev = torch.device("cuda:0")
torch.manual_seed(1234)
N, P = 100, 20
B = 148 * 2 * 4
env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
if os.environ.get("SLAP_FIXTURE_POLICY"):        # the reference-initialised fixture weights of the parity tests
    from tests.helpers import load_policy
    pol = load_policy("slap", dev)
else:
    pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
with torch.inference_mode():
    td = env.reset(env.generator(B).to(dev))
    h = pol.encoder(td)[0].float().contiguous()
W, q0, wcap, Wfixed = pol.decoder.folded_weights()
nt = W.shape[0] // 128
stream = _lib.stream_ptr(dev)
tables = torch.empty(nt, B * N, 128, device=dev)
gctx = torch.empty(B, 128, device=dev)
_lib.call("rl4co_am_project", B * N, nt, _lib.ptr(h), _lib.ptr(W), _lib.ptr(tables), stream)
_lib.call("rl4co_am_graph_context", B, N, _lib.ptr(h), _lib.ptr(Wfixed), _lib.ptr(gctx), stream)
trace = torch.zeros((P + 2) * 8, dtype=torch.int64, device=dev)
acts = torch.empty(B, P, dtype=torch.int64, device=dev)
logp = torch.empty(B, P, device=dev)
ll, rew = torch.empty(B, device=dev), torch.empty(B, device=dev)
steps = torch.empty(B, dtype=torch.int32, device=dev)
status = _lib.new_status(dev)
locs = td["locs"].contiguous()
pick = td["picklist"].contiguous()
a = _lib.AmRolloutArgs()
a.env, a.mode, a.B, a.S, a.multistart, a.num_loc = 2, _lib.MODE_GREEDY, B, 0, 0, N
a.N, a.T, a.P, a.O, a.K, a.n_tables = N, P, P, pick.shape[1], pick.shape[2], nt
a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
a.tables, a.gctx, a.locs = _lib.ptr(tables), _lib.ptr(gctx), _lib.ptr(locs)
a.picklist = _lib.ptr(pick, torch.int64)
a.actions, a.logp, a.logp_all, a.ll, a.reward = _lib.ptr(acts), _lib.ptr(logp), _lib.ptr(trace), _lib.ptr(ll), _lib.ptr(rew)
a.steps, a.status = _lib.ptr(steps), _lib.ptr(status)
for _ in range(2):
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
torch.cuda.synchronize()
rb = (steps.cpu().numpy() >> 8)
print("glimpse rebuilds per instance: mean %.2f  max %d  histogram %s" % (rb.mean(), rb.max(), np.bincount(rb)[:21].tolist()))
full = trace.cpu().numpy().reshape(P + 2, 8)
tr = full[:P, :5]
d = np.diff(tr, axis=1)
names = ["logits+butterfly", "tanh+warp argmax", "barrier", "winner+update"]
print("mean reward", rew.mean().item())
m = d[1:].mean(0)
print("per step: " + "  ".join(f"{n} {v:.0f}" for n, v in zip(names, m)), "| step total", (tr[1:, 0] - tr[:-1, 0]).mean())
print("cycles from step start to next step start:", " ".join(str(int(v)) for v in (tr[1:, 0] - tr[:-1, 0])))
print("update phase per step:", " ".join(str(int(v)) for v in d[:, 3]))
w = full[P + 1]
print(f"first glimpse state: V wait {w[6]-full[P][2]}  barrier {w[0]-w[6]}  exponentials + Z {w[1]-w[0]}  barrier {w[2]-w[1]}  "
      f"partial numerators {w[3]-w[2]}  barrier {w[4]-w[3]}  combine {w[5]-w[4]}")
v = full[P]
print(f"instance: K loads + scores {v[1]-v[0]}  small state {v[2]-v[1]}  V wait + first glimpse {v[3]-v[2]}  decode {v[4]-v[3]}  "
      f"log-probs {v[5]-v[4]}  reward + outputs {v[6]-v[5]}  total {v[6]-v[0]}")
