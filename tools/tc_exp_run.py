"""Time the tensor-core linear shapes with a given build of the library (tools/tc_experiments.py)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
from rl4co_slap_b200 import _lib
if len(sys.argv) > 1:
    _lib.LIB_PATH = os.path.abspath(sys.argv[1])
from rl4co_slap_b200.encoder import linear128
dev = torch.device("cuda:0")
M = 16384 * 100
def timed(fn, n=5):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
res = []
with torch.inference_mode():
    for n_in, n_out in ((128, 128), (128, 640), (512, 128)):
        lin = nn.Linear(n_in, n_out).to(dev)
        x = torch.randn(M, n_in, device=dev)
        res.append(f"{n_in}->{n_out} {timed(lambda: linear128(x, lin)):.3f} ms")
print(os.path.basename(os.path.dirname(_lib.LIB_PATH)), " | ".join(res))
