"""Encoder linears: torch fp32 F.linear (cuBLAS SIMT) vs rl4co_linear_tc, per shape; and attention."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn, torch.nn.functional as F
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
with torch.inference_mode():
    for n_in, n_out, relu in ((128, 384, False), (128, 128, False), (128, 512, True), (512, 128, False)):
        lin = nn.Linear(n_in, n_out).to(dev)
        x = torch.randn(M, n_in, device=dev)
        t_tc = timed(lambda: linear128(x, lin, relu))
        lin.tensor_core = False
        t_pt = timed(lambda: linear128(x, lin, relu))
        print(f"{n_in}->{n_out} relu={relu}: tc {t_tc:.3f} ms  torch {t_pt:.3f} ms")
    # skip + eval batch norm fused into the epilogue vs separate kernels
    for n_in in (128, 512):
        lin = nn.Linear(n_in, 128).to(dev)
        bn = nn.BatchNorm1d(128).to(dev).eval()
        x = torch.randn(M, n_in, device=dev)
        res = torch.randn(M, 128, device=dev)
        t_f = timed(lambda: linear128(x, lin, residual=res, norm=bn))
        t_s = timed(lambda: bn(linear128(x, lin) + res))
        t_r = timed(lambda: linear128(x, lin, residual=res))
        t_n = timed(lambda: linear128(x, lin, norm=bn))
        print(f"{n_in}->128 + skip + bn: fused {t_f:.3f} ms  separate {t_s:.3f} ms  (residual only {t_r:.3f}, affine only {t_n:.3f})")
