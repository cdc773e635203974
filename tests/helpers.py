"""Shared helpers of the parity tests (test infrastructure; may import oracle/)."""
from __future__ import annotations

import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TSP_CASES, CVRP_CASES, SLAP_CASES = ("tsp5", "tsp20", "tsp50", "tsp100"), ("cvrp10", "cvrp20", "cvrp50", "cvrp100"), ("slap50", "slap100")
OP_CASES = ("op10", "op20", "op50", "op100")
SDVRP_CASES = ("sdvrp10", "sdvrp20", "sdvrp50", "sdvrp100")
ATSP_CASES = ("atsp5", "atsp20", "atsp50", "atsp100")
PCTSP_CASES = ("pctsp10", "pctsp20", "pctsp50", "pctsp100", "spctsp20", "spctsp50")
_PCTSP_KEYS = ("locs", "depot", "penalty", "deterministic_prize", "stochastic_prize")


def fixture(name):
    return np.load(os.path.join(GOLDEN, f"{name}.npz"))


def env_of(case: str) -> str:
    return "".join(c for c in case.split("_")[0] if c.isalpha())


def t(x, dev=None):
    out = torch.from_numpy(np.asarray(x))
    return out.to(dev) if dev is not None else out


def policy_state(name: str):
    fx = fixture(f"policy_{name}")
    return {k: torch.from_numpy(fx[k]) for k in fx.files}


def load_policy(env_name: str, dev, variant: str = None):
    """The product policy carrying the fixture weights of the reference policy."""
    import rl4co_slap_b200 as R
    sd = policy_state(variant or env_name)
    kw = dict(env_name=env_name, num_encoder_layers=1, feedforward_hidden=128)
    if variant == "cvrp_pomo":
        kw.update(normalization="instance", use_graph_context=False)
    pol = R.AttentionModelPolicy(**kw)
    missing, unexpected = pol.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("num_batches_tracked" in m or "running_" in m for m in missing), missing
    return pol.to(dev).eval()


def oracle_weights(env_name: str, variant: str = None):
    sd = policy_state(variant or env_name)
    W = {"Wnode": sd["decoder.project_node_embeddings.weight"],
         "Wfixed": sd["decoder.project_fixed_context.weight"],
         "Wout": sd["decoder.pointer.project_out.weight"]}
    if env_name != "slap":
        W["Wctx"] = sd["decoder.context_embedding.project_context.weight"]
    if env_name == "tsp":
        W["Wph"] = sd["decoder.context_embedding.W_placeholder"]
    if env_name == "sdvrp":
        W["Wdyn"] = sd["decoder.dynamic_embedding.projection.weight"]
    return W


def gen_td(env_name: str, fx, dev=None):
    """The generator output stored in a fixture as a product TensorDict."""
    from rl4co_slap_b200 import TensorDict
    keys = {"tsp": ("locs",), "cvrp": ("locs", "depot", "demand", "capacity"),
            "slap": ("locs", "freq", "assignment", "picklist", "depot_loc_dist"),
            "op": ("locs", "depot", "prize", "max_length"), "pctsp": _PCTSP_KEYS, "spctsp": _PCTSP_KEYS,
            "atsp": ("cost_matrix",), "sdvrp": ("locs", "depot", "demand", "capacity")}[env_name]
    if env_name == "atsp":
        d = {"cost_matrix": t(fx["gen.cost_matrix"], dev)}
        return TensorDict(d, batch_size=[d["cost_matrix"].shape[0]])
    d = {k: t(fx[f"gen.{k}"], dev) for k in keys}
    return TensorDict(d, batch_size=[d["locs"].shape[0]])


def reset_from_fixture(env, env_name: str, fx, dev):
    return env.reset(gen_td(env_name, fx, dev))


def oracle_reset(env_name: str, fx):
    from oracle import port
    if env_name == "tsp":
        return port.tsp_reset(t(fx["gen.locs"]))
    if env_name == "cvrp":
        return port.cvrp_reset(t(fx["gen.depot"]), t(fx["gen.locs"]), t(fx["gen.demand"]),
                               float(fx["gen.vehicle_capacity"]))
    if env_name == "atsp":
        return port.atsp_reset(t(fx["gen.cost_matrix"]))
    if env_name == "sdvrp":
        return port.sdvrp_reset(t(fx["gen.depot"]), t(fx["gen.locs"]), t(fx["gen.demand"]),
                                float(fx["gen.vehicle_capacity"]))
    if env_name == "op":
        return port.op_reset(t(fx["gen.depot"]), t(fx["gen.locs"]), t(fx["gen.prize"]), t(fx["gen.max_length"]))
    if env_name in ("pctsp", "spctsp"):
        return port.pctsp_reset(*(t(fx[f"gen.{k}"]) for k in ("depot", "locs", "penalty", "deterministic_prize",
                                                               "stochastic_prize")),
                                prize_required=float(fx["gen.prize_required"]), stochastic=env_name == "spctsp")
    extra = {k: t(fx[f"gen.{k}"]) for k in ("freq", "picklist", "depot_loc_dist")}
    return port.slap_reset(t(fx["gen.locs"]), t(fx["gen.assignment"]), fx["gen.freq"].shape[-2], extra)
