"""Small workloads of the round-2 kernels for compute-sanitizer (memcheck / racecheck / synccheck), for machines where the
sanitizer may be used (it is closed on the pool this repository was built on, so no log of it is committed):
    compute-sanitizer --tool racecheck python tools/sanitize_new_kernels.py slap
    compute-sanitizer --tool memcheck  python tools/sanitize_new_kernels.py k3
slap: B = 700 instances (296 persistent CTAs, so most CTAs run more than two instances: the carry-over of registers /
shared memory between instances is exercised), greedy, sampling and evaluate; k3: 1 000 rows (ragged last tile) x 3 and 5 tables,
plus a 128 -> 512 -> 128 feed-forward pair (FP16 and TF32 forms, bias / ReLU / residual epilogues)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib

dev = torch.device("cuda:0")
which = sys.argv[1]
torch.manual_seed(0)
with torch.inference_mode():
    if which == "slap":
        env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
        pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
        td = env.reset(env.generator(700).to(dev))
        h = pol.encoder(td)[0]
        g = pol(td, env, decode_type="greedy", hidden=h, return_actions=True)
        s = pol(td, env, decode_type="sampling", hidden=h, return_actions=True)
        e = pol(td, env, actions=s["actions"], hidden=h)
        print("slap ok", float(g["reward"].mean()), float((e["log_likelihood"] - s["log_likelihood"]).abs().max()))
    else:
        st = _lib.stream_ptr(dev)
        for M, nt in ((1000, 3), (1000, 5), (128 * 9, 4)):
            hh = torch.randn(M, 128, device=dev) * 3
            W = torch.randn(nt * 128, 128, device=dev) * 0.1
            image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(nt)), device=dev)
            _lib.call("rl4co_am_pack_weights", nt, _lib.ptr(W), _lib.ptr(image), st)
            out = torch.zeros(nt, M, 128, device=dev)
            _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(hh), _lib.ptr(image), _lib.ptr(out), st)
            torch.cuda.synchronize()
            ref = (hh.double() @ W.double().t()).view(M, nt, 128).permute(1, 0, 2)
            print("k3", M, nt, float((out.double() - ref).abs().max()))
        from rl4co_slap_b200.encoder import linear128
        x = torch.randn(1000, 128, device=dev)
        l1, l2 = torch.nn.Linear(128, 512).to(dev), torch.nn.Linear(512, 128).to(dev)
        y = linear128(linear128(x, l1, relu=True), l2, residual=x)
        ref = torch.nn.functional.linear(torch.relu(torch.nn.functional.linear(x, l1.weight, l1.bias)), l2.weight, l2.bias) + x
        print("ffn", float((y - ref).abs().max()))
