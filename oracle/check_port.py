"""TEST INFRASTRUCTURE ONLY - pin oracle/port.py against the unmodified reference.

Runs ONLY in the build container (needs /root/reference via oracle/ref_shim.py):
    python oracle/check_port.py
For TSP / CVRP / SLAP it (1) replays random feasible action sequences through the
reference env and through the port, comparing every per-step state tensor bit for bit,
and (2) runs the reference AttentionModelPolicy and ``port.am_rollout`` on identical
weights / inputs for greedy, evaluate and multistart decoding, comparing tours exactly
and reward / log-likelihood to 1e-6.
"""
from __future__ import annotations

import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import port, ref_shim  # noqa: E402

ref_shim.install()
from rl4co.envs.routing.cvrp.env import CVRPEnv  # noqa: E402
from rl4co.envs.routing.tsp.env import TSPEnv  # noqa: E402
from rl4co.envs.warehousing.slap.env import SLAPEnv  # noqa: E402
from rl4co.models.zoo.am.policy import AttentionModelPolicy  # noqa: E402


def decoder_weights(policy, env_name):
    sd = policy.state_dict()
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"],
         "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"]}
    if env_name != "slap":
        W["Wctx"] = sd["decoder.context_embedding.project_context.weight"]
    if env_name == "tsp":
        W["Wph"] = sd["decoder.context_embedding.W_placeholder"]
    return W


def make(env_name, size, layers=1, normalization="batch", use_graph_context=True):
    if env_name == "tsp":
        env = TSPEnv(generator_params=dict(num_loc=size))
        pol = AttentionModelPolicy(env_name="tsp", num_encoder_layers=layers, normalization=normalization,
                                   use_graph_context=use_graph_context,
                                   context_embedding=ref_shim.upstream_tsp_context(128))
    elif env_name == "cvrp":
        env = CVRPEnv(generator_params=dict(num_loc=size))
        pol = AttentionModelPolicy(env_name="cvrp", num_encoder_layers=layers, normalization=normalization,
                                   use_graph_context=use_graph_context)
    else:
        env = SLAPEnv(generator_params={"n_aisles": size // 10, "n_locs": 10})
        ie, ce, de = ref_shim.slap_modules(128)
        pol = AttentionModelPolicy(env_name="slap", num_encoder_layers=layers, normalization=normalization,
                                   init_embedding=ie, context_embedding=ce, dynamic_embedding=de,
                                   use_graph_context=use_graph_context)
    return env, pol.eval()


def port_reset(env_name, td0, env):
    if env_name == "tsp":
        return port.tsp_reset(td0["locs"])
    if env_name == "cvrp":
        return port.cvrp_reset(td0["depot"], td0["locs"], td0["demand"], env.generator.vehicle_capacity)
    extra = {k: td0[k] for k in ("freq", "dist_mat", "picklist", "depot_loc_dist")}
    return port.slap_reset(td0["locs"], td0["assignment"], td0["freq"].shape[-2], extra)


def same(a, b, name):
    assert a.dtype == b.dtype, (name, a.dtype, b.dtype)
    assert a.shape == b.shape, (name, a.shape, b.shape)
    assert torch.equal(a, b), name


def check_replay(env_name, size, B):
    env, _ = make(env_name, size)
    td0 = env.generator(B)
    ref = env.reset(td0.clone())
    mine = port_reset(env_name, td0, env)
    keys = [k for k in ref.keys() if k in mine]
    for k in keys:
        same(ref[k], mine[k], f"{env_name} reset {k}")
    acts, steps = [], 0
    while not ref["done"].all():
        a = torch.multinomial(ref["action_mask"].float(), 1).squeeze(-1)
        ref.set("action", a)
        mine["action"] = a
        ref = env.step(ref)["next"]
        mine = port._STEP[env_name](mine)
        for k in [k for k in ref.keys() if k in mine]:
            same(ref[k], mine[k], f"{env_name} step {steps} {k}")
        acts.append(a)
        steps += 1
    acts = torch.stack(acts, 1)
    r_ref = env.get_reward(ref, acts)
    if env_name == "tsp":
        r = port.tsp_reward(mine["locs"], acts)
    elif env_name == "cvrp":
        port.cvrp_check(mine["demand"], mine["vehicle_capacity"], acts)
        r = port.cvrp_reward(mine["locs"], acts)
    else:
        r = port.slap_reward(mine["locs"], mine["assignment"], mine["picklist"])
    same(r_ref, r, f"{env_name} reward")
    print(f"  replay {env_name}{size} B={B}: {steps} steps, {len(keys)} keys bit-exact, reward bit-exact")


def check_policy(env_name, size, B, decode_type, **kw):
    env, pol = make(env_name, size, normalization=kw.pop("normalization", "batch"),
                    use_graph_context=kw.get("use_graph_context", True))
    td0 = env.generator(B)
    td = env.reset(td0.clone())
    W = decoder_weights(pol, env_name)
    with torch.inference_mode():
        ref_kw = {k: v for k, v in kw.items() if k != "use_graph_context"}
        out = pol(td.clone(), env, decode_type=decode_type, return_actions=True,
                  return_hidden=True, **ref_kw)
        h = out["hidden"].node_embeddings  # return_hidden yields the PrecomputedCache (base.py:226)
        mine = port.am_rollout(env_name, port_reset(env_name, td0, env), h, W, decode_type,
                               num_starts=kw.get("num_starts"),
                               use_graph_context=kw.get("use_graph_context", True),
                               num_loc=getattr(env.generator, "num_loc", None))
        same(out["actions"], mine["actions"], f"{env_name} {decode_type} actions")
        torch.testing.assert_close(out["reward"], mine["reward"], rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(out["log_likelihood"], mine["log_likelihood"], rtol=1e-6, atol=1e-5)
        # evaluate mode: given the reference's actions
        ev = pol(td.clone(), env, actions=out["actions"] if "multistart" not in decode_type else None,
                 decode_type=decode_type, return_actions=True) if "multistart" not in decode_type else None
        if ev is not None:
            mine_ev = port.am_rollout(env_name, port_reset(env_name, td0, env), h, W, actions=out["actions"])
            torch.testing.assert_close(ev["log_likelihood"], mine_ev["log_likelihood"], rtol=1e-6, atol=1e-5)
    print(f"  policy {env_name}{size} B={B} {decode_type} {kw}: tours exact, "
          f"T={out['actions'].shape[1]}, mean reward {out['reward'].mean():.4f}")


if __name__ == "__main__":
    torch.manual_seed(1234)
    for env_name, size in (("tsp", 20), ("tsp", 50), ("cvrp", 20), ("cvrp", 50), ("slap", 100)):
        check_replay(env_name, size, 16)
    for env_name, size in (("tsp", 20), ("tsp", 50), ("cvrp", 20), ("slap", 100)):
        check_policy(env_name, size, 8, "greedy")
    check_policy("tsp", 20, 4, "multistart_greedy", num_starts=20)
    check_policy("cvrp", 20, 4, "multistart_greedy", num_starts=20)
    check_policy("cvrp", 20, 4, "multistart_greedy", num_starts=20, normalization="instance",
                 use_graph_context=False)
    print("oracle/port.py matches the reference")
