"""TEST INFRASTRUCTURE ONLY - generate tests/golden/*.npz from the UNMODIFIED reference.

Runs only in the build container (imports /root/reference through oracle/ref_shim.py):
    python oracle/gen_golden.py
For each env it stores one small random-init policy (state_dict, 1 encoder layer, FFN 128)
and, per case, the generator output, a random feasible replay with every per-step state
tensor, and the reference policy's greedy / multistart-greedy tours, per-step log-probs,
rewards and log-likelihoods.  The fixtures are what pins oracle/port.py and the CUDA path
when /root/reference is not there (GPU box).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from rl4co.envs.routing.cvrp.env import CVRPEnv  # noqa: E402
from rl4co.envs.routing.tsp.env import TSPEnv  # noqa: E402
from rl4co.envs.warehousing.slap.env import SLAPEnv  # noqa: E402
from rl4co.models.zoo.am.policy import AttentionModelPolicy  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
GEN_KEYS = {"tsp": ("locs",), "cvrp": ("locs", "depot", "demand", "capacity"),
            "slap": ("locs", "freq", "assignment", "picklist", "depot_loc_dist")}


def np_(t):
    return t.detach().cpu().numpy()


def make_policy(env_name, **kw):
    common = dict(num_encoder_layers=1, feedforward_hidden=128, **kw)
    if env_name == "tsp":
        return AttentionModelPolicy(env_name="tsp", context_embedding=ref_shim.upstream_tsp_context(128),
                                    **common).eval()
    if env_name == "cvrp":
        return AttentionModelPolicy(env_name="cvrp", **common).eval()
    ie, ce, de = ref_shim.slap_modules(128)
    return AttentionModelPolicy(env_name="slap", init_embedding=ie, context_embedding=ce,
                                dynamic_embedding=de, **common).eval()


def make_env(env_name, size):
    if env_name == "tsp":
        return TSPEnv(generator_params=dict(num_loc=size))
    if env_name == "cvrp":
        return CVRPEnv(generator_params=dict(num_loc=size))
    return SLAPEnv(generator_params={"n_aisles": size // 10, "n_locs": 10})


def replay_trace(env, td0, weight=None):
    """Random feasible replay through the reference env; every state key after every step.
    ``weight`` optionally reshapes the sampling weights of the feasible actions."""
    td = env.reset(td0.clone())
    skip = set(GEN_KEYS[env.name]) | {"dist_mat", "terminated", "ratio"}
    trace = {f"reset.{k}": np_(v) for k, v in td.items() if k not in skip or k == "locs"}
    acts, t = [], 0
    while not td["done"].all():
        w = td["action_mask"].float()
        a = torch.multinomial(w if weight is None else weight(w), 1).squeeze(-1)
        td.set("action", a)
        td = env.step(td)["next"]
        for k, v in td.items():
            if k not in skip and k != "action":
                trace[f"step{t}.{k}"] = np_(v)
        acts.append(a)
        t += 1
    acts = torch.stack(acts, 1)
    trace["replay.actions"] = np_(acts)
    trace["replay.reward"] = np_(env.get_reward(td, acts))
    trace["replay.steps"] = np.int64(t)
    return trace


def policy_outputs(env, pol, td0, tag, **kw):
    td = env.reset(td0.clone())
    with torch.inference_mode():
        out = pol(td, env, return_actions=True, return_hidden=True,
                  return_sum_log_likelihood=False, **kw)
    res = {f"{tag}.actions": np_(out["actions"]), f"{tag}.reward": np_(out["reward"]),
           f"{tag}.logprobs": np_(out["log_likelihood"]).astype(np.float32),
           f"{tag}.ll": np_(out["log_likelihood"].sum(1)).astype(np.float32)}
    return res, out["hidden"].node_embeddings


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(1234)
    np.random.seed(1234)
    cases = {"tsp": [(5, 8), (20, 8), (50, 6), (100, 4)],
             "cvrp": [(10, 8), (20, 8), (50, 6), (100, 4)],
             "slap": [(100, 6), (50, 6)]}
    for env_name, lst in cases.items():
        pol = make_policy(env_name)
        np.savez_compressed(os.path.join(OUT, f"policy_{env_name}.npz"),
                            **{k: np_(v) for k, v in pol.state_dict().items()})
        for size, B in lst:
            env = make_env(env_name, size)
            td0 = env.generator(B)
            data = {f"gen.{k}": np_(td0[k]) for k in GEN_KEYS[env_name]}
            if env_name == "cvrp":
                data["gen.vehicle_capacity"] = np.float32(env.generator.vehicle_capacity)
            data.update(replay_trace(env, td0))
            res, h = policy_outputs(env, pol, td0, "greedy", decode_type="greedy")
            data.update(res)
            data["hidden"] = np_(h)
            if env_name != "slap":
                S = min(size, 8)
                res, _ = policy_outputs(env, pol, td0, "ms_greedy", decode_type="multistart_greedy",
                                        num_starts=S)
                data.update(res)
                data["ms_greedy.num_starts"] = np.int64(S)
            path = os.path.join(OUT, f"{env_name}{size}.npz")
            np.savez_compressed(path, **data)
            print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, greedy mean reward "
                  f"{data['greedy.reward'].mean():.4f}, T={data['greedy.actions'].shape[1]}")
    # POMO-style policy (instance norm, no graph context) for the CVRP multistart case
    pol = make_policy("cvrp", normalization="instance", use_graph_context=False)
    np.savez_compressed(os.path.join(OUT, "policy_cvrp_pomo.npz"),
                        **{k: np_(v) for k, v in pol.state_dict().items()})
    env = make_env("cvrp", 20)
    td0 = env.generator(6)
    data = {f"gen.{k}": np_(td0[k]) for k in GEN_KEYS["cvrp"]}
    data["gen.vehicle_capacity"] = np.float32(env.generator.vehicle_capacity)
    res, h = policy_outputs(env, pol, td0, "ms_greedy", decode_type="multistart_greedy", num_starts=20)
    data.update(res)
    data["hidden"] = np_(h)
    data["ms_greedy.num_starts"] = np.int64(20)
    np.savez_compressed(os.path.join(OUT, "cvrp20_pomo.npz"), **data)
    print("done")


if __name__ == "__main__":
    main()
