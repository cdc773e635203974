"""Side measurements of the other BASELINE.json configurations (parity-test cases, not bench lines):
C2 TSP100 env-only replay (batch 2^20), C4 CVRP100 POMO multistart sampling, C5 SLAP greedy sweep.
CUDA-event timed, 3 warm-ups, median of 5.  python tools/bench_configs.py > gpurun_out/configs.json"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rl4co_slap_b200 as R

dev = torch.device("cuda:0")
HBM = 6538.3e9


def timed(fn, warm=3, reps=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2] * 1e-3


out = {}
with torch.inference_mode():
    # ---- C2: TSP100 env-only replay, batch 2^20, fixed random permutations
    B, N = 1 << 20, 100
    torch.manual_seed(1234)
    env = R.TSPEnv(generator_params=dict(num_loc=N), device=dev, check_solution=True)
    td = env.reset(R.TensorDict({"locs": torch.rand(B, N, 2, device=dev)}, batch_size=[B]))
    acts = torch.rand(B, N, device=dev).argsort(1)
    s = timed(lambda: env.replay(td, acts))
    algo = 17 * N + 29
    out["C2_tsp100_replay"] = {"batch": B, "seconds": s, "rollouts_per_s": B / s, "env_steps_per_s": B * N / s,
                               "algorithmic_GBps": B * algo / s / 1e9, "frac_of_measured_hbm": B * algo / s / HBM}
    del td, acts
    # ---- C4: CVRP100 POMO multistart sampling, S = 100
    torch.manual_seed(1234)
    for Bc in (64, 256):
        env = R.CVRPEnv(generator_params=dict(num_loc=100), device=dev, check_solution=False)
        pol = R.POMOPolicy(env_name="cvrp").to(dev).eval()
        td = env.reset(env.generator(Bc).to(dev))
        h = pol.encoder(td)[0]
        s = timed(lambda: pol(td, env, decode_type="multistart_sampling", num_starts=100, hidden=h), warm=2, reps=3)
        o = pol(td, env, decode_type="multistart_sampling", num_starts=100, hidden=h, return_actions=True)
        T = o["actions"].shape[1]
        out[f"C4_cvrp100_pomo_B{Bc}"] = {"batch": Bc, "starts": 100, "steps": T, "seconds": s,
                                          "rollouts_per_s": Bc * 100 / s, "env_steps_per_s": Bc * 100 * T / s}
    # ---- C5: SLAP 10x10, 20 products, greedy sweep
    torch.manual_seed(1234)
    env = R.SLAPEnv(generator_params={"n_aisles": 10, "n_locs": 10}, device=dev)
    pol = R.AttentionModelPolicy(env_name="slap").to(dev).eval()
    td0 = env.reset(env.generator(1024).to(dev))
    for Bs in (64, 1024, 16384, 131072):
        rep = max(1, Bs // 1024)
        td = td0 if Bs == 1024 else (td0[:Bs] if Bs < 1024 else R.TensorDict(
            {k: v.repeat(rep, *([1] * (v.dim() - 1))) for k, v in td0.items()}, batch_size=[Bs]))
        h = torch.cat([pol.encoder(td[i:i + 8192])[0] for i in range(0, Bs, 8192)])
        s = timed(lambda: pol(td, env, decode_type="greedy", hidden=h), warm=2, reps=3)
        out[f"C5_slap_B{Bs}"] = {"batch": Bs, "seconds": s, "rollouts_per_s": Bs / s, "env_steps_per_s": Bs * 20 / s}
    # ---- sibling routing envs on the register-file kernel: 100 customers, greedy and sampling
    for name in ("op", "pctsp", "sdvrp"):
        torch.manual_seed(1234)
        Bs = 16384
        env = R.get_env(name, generator_params=dict(num_loc=100), device=dev, check_solution=False)
        pol = R.AttentionModelPolicy(env_name=name).to(dev).eval()
        td = env.reset(env.generator(Bs).to(dev))
        h = torch.cat([pol.encoder(td[i:i + 8192])[0] for i in range(0, Bs, 8192)])
        for mode in ("greedy", "sampling"):
            s = timed(lambda: pol(td, env, decode_type=mode, hidden=h), warm=2, reps=3)
            o = pol(td, env, decode_type=mode, hidden=h, return_actions=True)
            steps = float((o["actions"] != 0).sum(1).float().mean().item()) + 1
            out[f"{name}100_{mode}_B{Bs}"] = {"batch": Bs, "seconds": s, "rollouts_per_s": Bs / s, "mean_steps": steps,
                                              "env_steps_per_s": Bs * steps / s, "mean_reward": float(o["reward"].mean().item())}
print(json.dumps(out, indent=1))
