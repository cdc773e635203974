"""TEST INFRASTRUCTURE ONLY - tests/golden/op*.npz + policy_op.npz from the UNMODIFIED reference
(the orienteering problem, the sibling environment of SURVEY.md 8f rank 4).

Runs only in the build container (imports /root/reference through oracle/ref_shim.py):
    python oracle/gen_golden_op.py
Same content as oracle/gen_golden.py's cases (generator output, random feasible replay with every
per-step state tensor, greedy rollout of a small random-init reference policy) plus ``eval.*``: the
reference policy's per-step log-probs of the replay's random tours (decode_type="evaluate"), which
exercises long trajectories a random-init greedy policy does not produce.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import gen_golden as G  # noqa: E402  (installs the shim)

from rl4co.envs.routing.atsp.env import ATSPEnv  # noqa: E402
from rl4co.envs.routing.op.env import OPEnv  # noqa: E402
from rl4co.envs.routing.pctsp.env import PCTSPEnv  # noqa: E402
from rl4co.envs.routing.sdvrp.env import SDVRPEnv  # noqa: E402
from rl4co.envs.routing.spctsp.env import SPCTSPEnv  # noqa: E402
from rl4co.models.zoo.am.policy import AttentionModelPolicy  # noqa: E402

G.GEN_KEYS["op"] = ("locs", "depot", "prize", "max_length")
G.GEN_KEYS["pctsp"] = G.GEN_KEYS["spctsp"] = ("locs", "depot", "penalty", "deterministic_prize", "stochastic_prize")


def late_depot(w):
    """Return to the depot early only rarely, so that the replays are long tours that run into the length
    budget (a uniform draw over the feasible actions ends most tours within a few steps)."""
    w = w.clone()
    w[:, 0] = 1e-3
    return w


def main():
    torch.manual_seed(4321)
    np.random.seed(4321)
    pol = AttentionModelPolicy(env_name="op", num_encoder_layers=1, feedforward_hidden=128).eval()
    np.savez_compressed(os.path.join(G.OUT, "policy_op.npz"),
                        **{k: G.np_(v) for k, v in pol.state_dict().items()})
    for size, B in ((10, 8), (20, 8), (50, 6), (100, 4)):
        # the default prize_distribution trips torch's Uniform(1, 1) argument validation; "dist" is the
        # reference's own alternative (op/generator.py)
        env = OPEnv(generator_params=dict(num_loc=size, prize_distribution="dist"))
        td0 = env.generator(B)
        data = {f"gen.{k}": G.np_(td0[k]) for k in G.GEN_KEYS["op"]}
        data.update(G.replay_trace(env, td0, weight=late_depot))
        td_r = env.reset(td0.clone())       # reset-state tensors that share a name with a generator key
        data["reset.prize"], data["reset.max_length"] = G.np_(td_r["prize"]), G.np_(td_r["max_length"])
        res, h = G.policy_outputs(env, pol, td0, "greedy", decode_type="greedy")
        data.update(res)
        data["hidden"] = G.np_(h)
        res, _ = G.policy_outputs(env, pol, td0, "eval", actions=torch.from_numpy(data["replay.actions"]))
        data.update(res)
        S = min(size, 8)    # POMO starts; only where the reference does not resample them (utils/ops.py:166-176)
        if not bool((td_r["action_mask"][..., 1:].float().sum(-1) < S).any()):
            try:    # a forced start node beyond the budget fails the reference's own validity check: no fixture then
                res, _ = G.policy_outputs(env, pol, td0, "ms_greedy", decode_type="multistart_greedy", num_starts=S)
                data.update(res)
                data["ms_greedy.num_starts"] = np.int64(S)
            except AssertionError as e:
                print(f"op{size}: reference multistart rejected its own tour ({e})")
        path = os.path.join(G.OUT, f"op{size}.npz")
        np.savez_compressed(path, **data)
        print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, greedy mean reward "
              f"{data['greedy.reward'].mean():.4f}, T={data['greedy.actions'].shape[1]}, "
              f"replay T={data['replay.actions'].shape[1]}")


def main_pctsp():
    """pctsp / spctsp (same policy class; spctsp collects the stochastic prize)."""
    torch.manual_seed(8765)
    np.random.seed(8765)
    for name, cls in (("pctsp", PCTSPEnv), ("spctsp", SPCTSPEnv)):
        pol = AttentionModelPolicy(env_name=name, num_encoder_layers=1, feedforward_hidden=128).eval()
        np.savez_compressed(os.path.join(G.OUT, f"policy_{name}.npz"),
                            **{k: G.np_(v) for k, v in pol.state_dict().items()})
        for size, B in (((10, 8), (20, 8), (50, 6), (100, 4)) if name == "pctsp" else ((20, 8), (50, 6))):
            env = cls(generator_params=dict(num_loc=size))
            td0 = env.generator(B)
            data = {f"gen.{k}": G.np_(td0[k]) for k in G.GEN_KEYS[name]}
            data["gen.prize_required"] = np.float32(env.generator.prize_required)
            data.update(G.replay_trace(env, td0))
            td_r = env.reset(td0.clone())
            data["reset.penalty"] = G.np_(td_r["penalty"])
            res, h = G.policy_outputs(env, pol, td0, "greedy", decode_type="greedy")
            data.update(res)
            data["hidden"] = G.np_(h)
            res, _ = G.policy_outputs(env, pol, td0, "eval", actions=torch.from_numpy(data["replay.actions"]))
            data.update(res)
            S = min(size, 8)
            res, _ = G.policy_outputs(env, pol, td0, "ms_greedy", decode_type="multistart_greedy", num_starts=S)
            data.update(res)
            data["ms_greedy.num_starts"] = np.int64(S)
            path = os.path.join(G.OUT, f"{name}{size}.npz")
            np.savez_compressed(path, **data)
            print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, greedy mean reward "
                  f"{data['greedy.reward'].mean():.4f}, T={data['greedy.actions'].shape[1]}, "
                  f"replay T={data['replay.actions'].shape[1]}")


def main_atsp():
    """atsp: environment only (generator incl. the triangle-inequality closure, replay states, reward)."""
    torch.manual_seed(2468)
    G.GEN_KEYS["atsp"] = ()
    for size, B in ((5, 8), (20, 8), (50, 6), (100, 4)):
        env = ATSPEnv(generator_params=dict(num_loc=size))
        td0 = env.generator(B)
        data = {"gen.cost_matrix": G.np_(td0["cost_matrix"])}
        tr = G.replay_trace(env, td0)
        data.update({k: v for k, v in tr.items() if not k.endswith(".cost_matrix")})
        path = os.path.join(G.OUT, f"atsp{size}.npz")
        np.savez_compressed(path, **data)
        print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, replay mean reward {data['replay.reward'].mean():.4f}")


def main_sdvrp():
    """sdvrp: split deliveries (nodes revisited) and the per-step dynamic embedding of the remaining demands."""
    torch.manual_seed(1357)
    G.GEN_KEYS["sdvrp"] = G.GEN_KEYS["cvrp"]
    pol = AttentionModelPolicy(env_name="sdvrp", num_encoder_layers=1, feedforward_hidden=128).eval()
    np.savez_compressed(os.path.join(G.OUT, "policy_sdvrp.npz"), **{k: G.np_(v) for k, v in pol.state_dict().items()})
    for size, B in ((10, 8), (20, 8), (50, 6), (100, 4)):
        env = SDVRPEnv(generator_params=dict(num_loc=size))
        td0 = env.generator(B)
        data = {f"gen.{k}": G.np_(td0[k]) for k in G.GEN_KEYS["sdvrp"]}
        data["gen.vehicle_capacity"] = np.float32(env.generator.vehicle_capacity)
        data.update(G.replay_trace(env, td0))
        res, h = G.policy_outputs(env, pol, td0, "greedy", decode_type="greedy")
        data.update(res)
        data["hidden"] = G.np_(h)
        res, _ = G.policy_outputs(env, pol, td0, "eval", actions=torch.from_numpy(data["replay.actions"]))
        data.update(res)
        path = os.path.join(G.OUT, f"sdvrp{size}.npz")
        np.savez_compressed(path, **data)
        print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, greedy mean reward {data['greedy.reward'].mean():.4f}, "
              f"T={data['greedy.actions'].shape[1]}, replay T={data['replay.actions'].shape[1]}")


if __name__ == "__main__":
    if "--sdvrp-only" in sys.argv:
        main_sdvrp()
        sys.exit(0)
    if "--atsp-only" in sys.argv:
        main_atsp()
        sys.exit(0)
    if "--pctsp-only" not in sys.argv:
        main()
    main_pctsp()
    main_atsp()
    main_sdvrp()
