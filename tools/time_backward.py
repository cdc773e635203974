"""reinforce_step on TSP100 / CVRP100: the backward kernel (csrc/am_backward.cu) beside autograd through the batched
PyTorch pass - time and peak memory of one training step (rollout + differentiable re-evaluation + backward).
    python tools/time_backward.py [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R
from rl4co_slap_b200.training import kernel_log_likelihood, reinforce_step, teacher_forced_log_likelihood

dev = torch.device("cuda:0")
cases = [("tsp", "sampling", None, int(b)) for b in (sys.argv[1:] or ["2048", "16384"])] + [("cvrp", "multistart_sampling", 100, 64)]
for name, dt, S, B in cases:
    torch.manual_seed(1234)
    pol = (R.POMOPolicy(env_name=name) if S else R.AttentionModelPolicy(env_name=name)).to(dev)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=100), check_solution=False)
    td = env.reset(env.generator(B).to(dev))
    opt = torch.optim.Adam(pol.parameters(), lr=1e-4)
    for mode in ("kernel", "torch"):
        try:
            ts = []
            torch.cuda.reset_peak_memory_stats()
            for it in range(4):
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                out = reinforce_step(pol, env, td, opt, baseline="shared" if S else None, decode_type=dt, num_starts=S, backward=mode)
                e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            print(f"{name}100 B={B} S={S} backward={mode:6s}: {sorted(ts[1:])[1]:9.1f} ms per training step, peak memory "
                  f"{torch.cuda.max_memory_allocated() / 2**30:.2f} GiB, loss {out['loss'].item():.4f}", flush=True)
        except torch.OutOfMemoryError as e:
            print(f"{name}100 B={B} S={S} backward={mode:6s}: out of memory ({str(e)[:80]})", flush=True)
            torch.cuda.empty_cache()

    # ---- the decoder part alone: log-likelihood forward + backward of given tours and a given encoder output
    with torch.no_grad():
        h0 = pol.encoder(td)[0].float()
        kw = dict(decode_type=dt, num_starts=S) if S else dict(decode_type=dt)
        acts = pol(td, env, return_actions=True, hidden=h0, **kw)["actions"]
    wts = torch.randn(acts.shape[0], device=dev)
    for mode, fn in (("kernel", kernel_log_likelihood), ("torch", teacher_forced_log_likelihood)):
        try:
            ts = []
            torch.cuda.reset_peak_memory_stats()
            for it in range(4):
                h = h0.clone().requires_grad_(True)
                pol.zero_grad(set_to_none=True)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                (wts * fn(pol, td, name, h, acts, num_starts=S or 0)).sum().backward()
                e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            print(f"   decoder only (ll forward + backward): {mode:6s} {sorted(ts[1:])[1]:9.1f} ms, peak memory "
                  f"{torch.cuda.max_memory_allocated() / 2**30:.2f} GiB", flush=True)
        except torch.OutOfMemoryError:
            print(f"   decoder only: {mode:6s} out of memory", flush=True)
            torch.cuda.empty_cache()
