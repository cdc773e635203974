"""Encoder layer timing: fused skip + batch-norm epilogue vs separate PyTorch ops (eval mode)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rl4co_slap_b200.encoder import GraphAttentionNetwork
dev = torch.device("cuda:0")
B, N = 16384, 100
def timed(fn, n=3):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
with torch.inference_mode():
    net = GraphAttentionNetwork(8, 128, 3).to(dev).eval()
    x = torch.randn(B, N, 128, device=dev)
    t_f = timed(lambda: net(x))
    for l in net.layers: l.fused_epilogue = False
    t_u = timed(lambda: net(x))
    print(f"3 layers, {B} x {N}: fused epilogue {t_f:.2f} ms, separate add + batch norm {t_u:.2f} ms")
