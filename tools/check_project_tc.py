"""GPU check of the tensor-core projection (csrc/am_project_tc.cu) against fp64 and the fp32 SIMT kernel."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rl4co_slap_b200 import _lib
dev = torch.device("cuda:0")
torch.manual_seed(0)
for M, nt in ((128, 1), (300, 3), (16384 * 100 // 16, 5), (131072 * 100 // 8, 5)):
    h = torch.randn(M, 128, device=dev) * 3
    W = torch.randn(nt * 128, 128, device=dev) * 0.1
    s = _lib.stream_ptr(dev)
    image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(nt)), device=dev)
    _lib.call("rl4co_am_pack_weights", nt, _lib.ptr(W), _lib.ptr(image), s)
    out_tc = torch.zeros(nt, M, 128, device=dev)
    out_simt = torch.zeros(nt, M, 128, device=dev)
    print("launch", M, nt, flush=True)
    _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(h), _lib.ptr(image), _lib.ptr(out_tc), s)
    torch.cuda.synchronize()
    _lib.call("rl4co_am_project", M, nt, _lib.ptr(h), _lib.ptr(W), _lib.ptr(out_simt), s)
    torch.cuda.synchronize()
    rows = slice(0, min(M, 4096))
    ref = (h[rows].double() @ W.double().t()).view(-1, nt, 128).permute(1, 0, 2)
    scale = ref.abs().max().item()
    e_tc = (out_tc[:, rows].double() - ref).abs().max().item()
    e_simt = (out_simt[:, rows].double() - ref).abs().max().item()
    d = (out_tc - out_simt).abs().max().item()
    print(f"  max|ref| {scale:.3f}  err tc {e_tc:.3e}  err simt {e_simt:.3e}  tc-simt (all rows) {d:.3e}", flush=True)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    for _ in range(3):
        _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(h), _lib.ptr(image), _lib.ptr(out_tc), s)
    ev[1].record()
    for _ in range(3):
        _lib.call("rl4co_am_project", M, nt, _lib.ptr(h), _lib.ptr(W), _lib.ptr(out_simt), s)
    ev[2].record()
    torch.cuda.synchronize()
    t_tc, t_simt = ev[0].elapsed_time(ev[1]) / 3, ev[1].elapsed_time(ev[2]) / 3
    gb = (M * 512 * (1 + nt)) / 1e9
    print(f"  tc {t_tc:.3f} ms ({gb / t_tc * 1e3:.0f} GB/s)  simt {t_simt:.3f} ms", flush=True)
