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


# ---- tolerances of the parity tests ------------------------------------------------------------------------------
# Summed log-likelihoods and rewards: 1e-5 relative (BASELINE.json north_star).  A PER-STEP log-prob is
# logit_a - logsumexp(logits) with |logit| <= tanh_clipping = 10: the fp32 resolution of the logits themselves
# (ulp(10) = 9.5e-7, a sum over <= 128 exponentials, the 3xTF32 table projection's ~2e-5 absolute error on entries of
# size ~18 feeding a 128-term dot product) is an ABSOLUTE floor of ~1e-5 whatever the log-prob's size (a log-prob of
# -1e-3 cannot be held to 1e-8), so per-step values are held to 1e-5 relative + 2e-5 absolute.
LOGP_RTOL, LOGP_ATOL = 1e-5, 2e-5
NEAR_TIE_EPS = 1e-4     # two candidates whose fp32 log-probs differ by less than 10 ulp of the clip bound are a tie


def assert_logp_close(got, want, noise=None):
    """|got - want| <= LOGP_RTOL |want| + LOGP_ATOL (+ 4 x the reference's OWN fp32 rounding noise when given).
    ``noise`` [B,T] = |fp32 reference - fp64 run of the same reference| per step; the largest value of an instance's
    trajectory is that instance's conditioning (tsp100 / cvrp100 fixtures: 5e-7, slap100 with |h| up to 33: 2.2e-5) -
    nobody can be held closer to an fp32 reference than that reference is to the exact result."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    err, bound = np.abs(got - want), LOGP_RTOL * np.abs(want) + LOGP_ATOL
    if noise is not None:
        bound = bound + 4.0 * np.asarray(noise, dtype=np.float64).max(axis=-1, keepdims=True)
    assert (err <= bound).all(), f"per-step log-prob off by {err.max():.3e} (worst excess {(err - bound).max():.3e})"


def assert_divergences_are_near_ties(env_name, state, h, W, a, b, eps=NEAR_TIE_EPS, **rollout_kw):
    """``a`` and ``b`` [B,T] are two greedy tour sets for the same instances.  For every instance where they differ,
    take the FIRST differing step t (the prefixes agree, so both kernels saw the same state there), evaluate the
    oracle's log-probs at that step on the common prefix and require that the two chosen nodes are tied to within
    ``eps`` AND are the oracle's two best candidates (i.e. the flip is an fp32 rounding artefact, not a bug).
    Returns (number of diverging instances, largest gap seen)."""
    from oracle import port
    a, b = torch.as_tensor(a).cpu(), torch.as_tensor(b).cpu()
    T = max(a.shape[1], b.shape[1])
    a = torch.nn.functional.pad(a, (0, T - a.shape[1]))
    b = torch.nn.functional.pad(b, (0, T - b.shape[1]))
    rows = (a != b).any(1).nonzero().flatten()
    if rows.numel() == 0:
        return 0, 0.0
    first = (a[rows] != b[rows]).float().argmax(1)
    sub = {k: (v[rows] if isinstance(v, torch.Tensor) and v.dim() > 0 and v.shape[0] == a.shape[0] else v)
           for k, v in state.items()}
    ref = port.am_rollout(env_name, sub, h[rows].cpu(), W, actions=a[rows], return_all_logp=True,
                          check_solution=False, **rollout_kw)
    lp = ref["logprobs"]                                    # [rows, T, N]
    worst = 0.0
    for j in range(rows.numel()):
        t0 = int(first[j])
        row = lp[j, t0]
        ca, cb = int(a[rows[j], t0]), int(b[rows[j], t0])
        gap = abs(float(row[ca]) - float(row[cb]))
        top = float(row.max())
        assert gap < eps, f"instance {int(rows[j])} step {t0}: chosen nodes {ca}/{cb} differ by {gap:.3e} in log-prob"
        assert top - min(float(row[ca]), float(row[cb])) < eps, \
            f"instance {int(rows[j])} step {t0}: a chosen node is {top - min(float(row[ca]), float(row[cb])):.3e} below the best"
        worst = max(worst, gap)
    return int(rows.numel()), worst
