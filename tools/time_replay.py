"""C2: TSP100 env-only replay, B = 2^20: CUDA-event time of rl4co_tsp_replay and its fraction of the measured HBM peak."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib
dev = torch.device("cuda:0")
B, N = 1 << 20, 100
torch.manual_seed(1234)
locs = torch.rand(B, N, 2, device=dev)
acts = torch.rand(B, N, device=dev).argsort(1)
o = {"first": torch.empty(B, dtype=torch.int64, device=dev), "cur": torch.empty(B, dtype=torch.int64, device=dev),
     "i": torch.empty(B, 1, dtype=torch.int64, device=dev), "mask": torch.empty(B, N, dtype=torch.uint8, device=dev),
     "done": torch.empty(B, dtype=torch.uint8, device=dev), "reward": torch.empty(B, device=dev)}
status = _lib.new_status(dev)
s = _lib.stream_ptr(dev)
def run():
    _lib.call("rl4co_tsp_replay", B, N, N, _lib.ptr(locs), _lib.ptr(acts), _lib.ptr(o["first"]), _lib.ptr(o["cur"]), _lib.ptr(o["i"]),
              _lib.ptr(o["mask"]), _lib.ptr(o["done"]), _lib.ptr(o["reward"]), 1, _lib.ptr(status), s)
for _ in range(5): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 200
try: peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception: peak = 6650.0
gb = B * 1729 / (ms * 1e-3) / 1e9
# correctness against torch on a sub-sample
idx = torch.arange(0, B, B // 4096, device=dev)
p = locs[idx].gather(1, acts[idx][..., None].expand(-1, -1, 2))
ref = -(p.roll(-1, 1) - p).norm(dim=-1).sum(1)
print(f"tsp_replay B=2^20: {ms:.4f} ms  {B / ms * 1e-6:.3f} G rollouts/s  {gb:.0f} GB/s algorithmic = {gb / peak:.3f} of measured HBM; "
      f"status {int(status.item())}, all done {bool(o['done'].all())}, mask zero {bool((o['mask'] == 0).all())}, "
      f"max rel reward err {((o['reward'][idx] - ref).abs() / ref.abs()).max().item():.2e}")
