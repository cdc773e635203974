"""Workloads of the side configurations for ncu captures (one kernel each, run under `ncu -k regex:...`):
    python tools/profile_side.py c2      TSP100 env-only replay, B = 2^20           (kernel: tsp_replay)
    python tools/profile_side.py c5      SLAP greedy rollout, B = 16 384            (kernel: am_rollout_slap)
    python tools/profile_side.py k3 [nt] K3 projection, 1 638 400 rows x nt tables  (kernel: am_project_tc)
    python tools/profile_side.py c4      CVRP100 POMO multistart sampling, B = 64   (kernel: the multistart rollout)
    python tools/profile_side.py attention  encoder self-attention, 16 384 x 100 nodes (kernel: encoder_attention_tc3)
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R

dev = torch.device("cuda:0")
which = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 and which != "k3" else 3
with torch.inference_mode():
    torch.manual_seed(1234)
    if which == "c2":
        B, N = 1 << 20, 100
        env = R.TSPEnv(generator_params=dict(num_loc=N), device=dev, check_solution=True)
        td = R.TensorDict({"locs": torch.rand(B, N, 2, device=dev)}, batch_size=[B])
        acts = torch.rand(B, N, device=dev).argsort(1)
        for _ in range(reps):
            env.replay(td, acts)
    elif which == "c5":
        Bs = 16384
        env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
        pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
        td = env.reset(env.generator(Bs).to(dev))
        h = torch.cat([pol.encoder(td[i:i + 8192])[0] for i in range(0, Bs, 8192)])
        for _ in range(reps):
            pol(td, env, decode_type="greedy", hidden=h)
    elif which == "k3":
        from rl4co_slap_b200 import _lib
        M, nt = 1638400, int(sys.argv[2]) if len(sys.argv) > 2 else 5
        h = torch.randn(M, 128, device=dev) * 3
        W = torch.randn(nt * 128, 128, device=dev) * 0.1
        s = _lib.stream_ptr(dev)
        image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(nt)), device=dev)
        _lib.call("rl4co_am_pack_weights", nt, _lib.ptr(W), _lib.ptr(image), s)
        out = torch.zeros(nt, M, 128, device=dev)
        for _ in range(3):
            _lib.call("rl4co_am_project_tc", M, nt, _lib.ptr(h), _lib.ptr(image), _lib.ptr(out), s)
    elif which == "c4":
        Bc = 64
        env = R.CVRPEnv(generator_params=dict(num_loc=100), device=dev, check_solution=False)
        pol = R.POMOPolicy(env_name="cvrp").to(dev).eval()
        td = env.reset(env.generator(Bc).to(dev))
        h = pol.encoder(td)[0]
        for _ in range(reps):
            pol(td, env, decode_type="multistart_sampling", num_starts=100, hidden=h)
    elif which == "attention":
        from rl4co_slap_b200 import _lib
        B, N = 16384, 100
        qkv = torch.randn(B, N, 384, device=dev)
        out = torch.empty(B, N, 128, device=dev)
        for _ in range(3):
            _lib.call("rl4co_encoder_attention_tc", B, N, _lib.ptr(qkv), _lib.ptr(out), _lib.stream_ptr(dev))
    torch.cuda.synchronize()
print("done", which)
