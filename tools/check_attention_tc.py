"""Encoder attention: tensor-core Q K^T kernel vs the SIMT kernel vs an fp64 reference; timing."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rl4co_slap_b200 import _lib
dev = torch.device("cuda:0")
def ref64(qkv, B, N):
    q, k, v = qkv.double().view(B, N, 3, 8, 16).permute(2, 0, 3, 1, 4).unbind(0)
    a = torch.softmax(q @ k.transpose(-1, -2) * 0.25, -1)
    return (a @ v).transpose(1, 2).reshape(B, N, 128)
def run(name, B, N, qkv):
    out = torch.empty(B, N, 128, device=dev)
    _lib.call(name, B, N, _lib.ptr(qkv), _lib.ptr(out), _lib.stream_ptr(dev))
    torch.cuda.synchronize()
    return out
def timed(fn, n=5):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
torch.manual_seed(0)
for B, N, amp in ((3, 100, 1.0), (7, 5, 1.0), (300, 50, 3.0), (150, 128, 1.0), (1000, 101, 2.0), (2, 33, 6.0)):
    qkv = torch.randn(B, N, 384, device=dev) * amp
    r = ref64(qkv, B, N)
    _lib.call("rl4co_set_attention_impl", 1)
    tc = run("rl4co_encoder_attention_tc", B, N, qkv)
    e2 = e3 = e4 = float("nan")
    if N <= 112:
        _lib.call("rl4co_set_attention_impl", 2)
        tc2 = run("rl4co_encoder_attention_tc", B, N, qkv)
        e2 = (tc2.double() - r).abs().max().item()
        _lib.call("rl4co_set_attention_impl", 3)
        tc3 = run("rl4co_encoder_attention_tc", B, N, qkv)
        e3 = (tc3.double() - r).abs().max().item()
        _lib.call("rl4co_set_attention_impl", 4)
        tc4 = run("rl4co_encoder_attention_tc", B, N, qkv)
        e4 = (tc4.double() - r).abs().max().item()
    si = run("rl4co_encoder_attention", B, N, qkv)
    s = r.abs().max().item()
    print(f"B={B} N={N} amp={amp}: max|ref| {s:.3f}  err tc(QK) {(tc.double()-r).abs().max().item():.3e}  err tc(QK+PV) {e2:.3e}  err tc(QK + fp16 PV) {e3:.3e} / 4 groups {e4:.3e}  err simt {(si.double()-r).abs().max().item():.3e}", flush=True)
B, N = 16384, 100
qkv = torch.randn(B, N, 384, device=dev)
out = torch.empty(B, N, 128, device=dev)
for name, impl in (("rl4co_encoder_attention_tc", 4), ("rl4co_encoder_attention_tc", 3), ("rl4co_encoder_attention_tc", 2), ("rl4co_encoder_attention_tc", 1), ("rl4co_encoder_attention", 0)):
    _lib.call("rl4co_set_attention_impl", impl)
    t = timed(lambda: _lib.call(name, B, N, _lib.ptr(qkv), _lib.ptr(out), _lib.stream_ptr(dev)))
    print(f"{name} impl {impl}: {t:.3f} ms per {B} x {N}", flush=True)
