"""TEST INFRASTRUCTURE ONLY - CPU restatement ("port") of the reference rollout path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this module; the product (``rl4co_slap_b200``) never does and
fails loudly when its CUDA library is missing.

Parity status: PINNED against the unmodified reference run through ``oracle/ref_shim.py``
(``oracle/check_port.py`` here; committed fixtures ``tests/golden/*.npz`` produced by
``oracle/gen_golden.py``).  The reference's own tests hold no numeric golden for this path
(SURVEY.md 8c), so the fixtures generated from the reference code are the anchor.
torchrl's ``EnvBase.reset`` merge behaviour is restated, not executed (unpinned third
party, see ref_shim.py header).

Everything is plain torch-CPU fp32 / int64 tensor arithmetic in the reference's op order,
state is a python dict of tensors.  Each function cites the reference file:line it follows
(paths relative to /root/reference/rl4co).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

State = Dict[str, torch.Tensor]


# =====================================================================================
# helpers  (utils/ops.py)
# =====================================================================================
def gather_by_index(src, idx, dim=1, squeeze=True):
    """utils/ops.py:65-77."""
    shape = list(src.shape)
    shape[dim] = -1
    idx = idx.view(idx.shape + (1,) * (src.dim() - idx.dim())).expand(shape)
    out = src.gather(dim, idx)
    return out.squeeze(dim) if (idx.size(dim) == 1 and squeeze) else out


def tour_length(ordered_locs):
    """utils/ops.py:88-101: closed tour, L2 norm of consecutive differences, summed."""
    nxt = torch.roll(ordered_locs, -1, dims=-2)
    return (nxt - ordered_locs).norm(p=2, dim=-1).sum(-1)


def batchify(x, repeats: int):
    """utils/ops.py:11-34, start-major: out[s*B + b] = x[b]."""
    if repeats <= 0:
        return x
    if isinstance(x, dict):
        return {k: batchify(v, repeats) for k, v in x.items()}
    s = x.shape
    return x.expand(repeats, *s).contiguous().view(s[0] * repeats, *s[1:])


def unbatchify(x, repeats: int):
    """utils/ops.py:37-62: [(S B), ...] -> [B, S, ...]."""
    if repeats <= 0:
        return x
    if isinstance(x, dict):
        return {k: unbatchify(v, repeats) for k, v in x.items()}
    s = x.shape
    return x.view(repeats, s[0] // repeats, *s[1:]).permute(1, 0, *range(2, len(s) + 1))


def num_starts_for(env_name: str, action_mask) -> int:
    """utils/ops.py:126-136."""
    n = action_mask.shape[-1]
    return n - 1 if env_name in ("cvrp", "sdvrp", "op", "pctsp", "spctsp") else n


def select_start_nodes(env_name: str, batch: int, num_starts: int, num_loc: int):
    """utils/ops.py:139-163 (tsp / depot branch)."""
    sel = torch.arange(num_starts).repeat_interleave(batch) % num_loc
    return sel if env_name == "tsp" else sel + 1


# =====================================================================================
# TSP  (envs/routing/tsp/env.py)
# =====================================================================================
def tsp_reset(locs) -> State:
    """tsp/env.py:95-120 + EnvBase.reset merge (done/terminated [B,1] False)."""
    B, N, _ = locs.shape
    cur = torch.zeros(B, dtype=torch.int64)
    return {
        "locs": locs,
        "first_node": cur,
        "current_node": cur,
        "i": torch.zeros(B, 1, dtype=torch.int64),
        "action_mask": torch.ones(B, N, dtype=torch.bool),
        "reward": torch.zeros(B, 1, dtype=torch.float32),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }


def tsp_step(td: State) -> State:
    """tsp/env.py:67-93.  Note the batch-global first-step test (`td["i"].all() == 0`,
    :70), done of shape [B], reward = zeros_like(done) (bool)."""
    cur = td["action"]
    first = cur if bool(td["i"].all() == 0) else td["first_node"]
    avail = td["action_mask"].scatter(-1, cur.unsqueeze(-1).expand_as(td["action_mask"]), 0)
    done = torch.sum(avail, dim=-1) == 0
    td.update({
        "first_node": first, "current_node": cur, "i": td["i"] + 1,
        "action_mask": avail, "reward": torch.zeros_like(done), "done": done,
    })
    return td


def tsp_check(actions) -> None:
    """tsp/env.py:165-173."""
    N = actions.size(1)
    ok = (torch.arange(N).view(1, -1).expand_as(actions) == actions.sort(1)[0]).all()
    assert ok, "Invalid tour"


def tsp_reward(locs, actions, check: bool = True):
    """tsp/env.py:157-163."""
    if check:
        tsp_check(actions)
    return -tour_length(gather_by_index(locs, actions))


# =====================================================================================
# CVRP  (envs/routing/cvrp/env.py)
# =====================================================================================
def cvrp_action_mask(td: State):
    """cvrp/env.py:137-149."""
    exceeds = td["demand"] + td["used_capacity"] > td["vehicle_capacity"]
    mask_loc = td["visited"][..., 1:].to(exceeds.dtype) | exceeds
    mask_depot = (td["current_node"] == 0) & ((mask_loc == 0).int().sum(-1) > 0)[:, None]
    return ~torch.cat((mask_depot, mask_loc), -1)


def cvrp_reset(depot, locs, demand, vehicle_capacity: float = 1.0) -> State:
    """cvrp/env.py:107-135 (+ merge).  ``demand`` is already divided by capacity
    (cvrp/generator.py:135-142)."""
    B, n, _ = locs.shape
    td = {
        "locs": torch.cat((depot[:, None, :], locs), -2),
        "demand": demand,
        "current_node": torch.zeros(B, 1, dtype=torch.int64),
        "used_capacity": torch.zeros(B, 1),
        "vehicle_capacity": torch.full((B, 1), vehicle_capacity),
        "visited": torch.zeros(B, n + 1, dtype=torch.uint8),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }
    td["action_mask"] = cvrp_action_mask(td)
    return td


def cvrp_step(td: State) -> State:
    """cvrp/env.py:73-105."""
    cur = td["action"][:, None]
    n_loc = td["demand"].size(-1)
    sel = gather_by_index(td["demand"], torch.clamp(cur - 1, 0, n_loc - 1), squeeze=False)
    used = (td["used_capacity"] + sel) * (cur != 0).float()
    visited = td["visited"].scatter(-1, cur, 1)
    done = visited.sum(-1) == visited.size(-1)
    td.update({"current_node": cur, "used_capacity": used, "visited": visited,
               "reward": torch.zeros_like(done), "done": done})
    td["action_mask"] = cvrp_action_mask(td)
    return td


def cvrp_check(demand, vehicle_capacity, actions) -> None:
    """cvrp/env.py:162-190."""
    B, n = demand.size()
    sorted_pi = actions.sort(1)[0]
    ok = (torch.arange(1, n + 1).view(1, -1).expand(B, n) == sorted_pi[:, -n:]).all() and (
        sorted_pi[:, :-n] == 0).all()
    assert ok, "Invalid tour"
    d = torch.cat((-vehicle_capacity, demand), 1).gather(1, actions)
    used = torch.zeros_like(demand[:, 0])
    for t in range(actions.size(1)):
        used += d[:, t]
        used[used < 0] = 0
        assert (used <= vehicle_capacity[:, 0] + 1e-5).all(), "Used more than capacity"


def cvrp_reward(locs_with_depot, actions):
    """cvrp/env.py:151-160 (the validity check is called from get_reward, base.py:182-188)."""
    ordered = torch.cat([locs_with_depot[..., 0:1, :], gather_by_index(locs_with_depot, actions)], 1)
    return -tour_length(ordered)


# =====================================================================================
# OP  (envs/routing/op/env.py) - sibling env of SURVEY.md 8f rank 4
# =====================================================================================
def op_action_mask(td: State):
    """op/env.py:154-170: visited, depot already revisited, or arriving would exceed the per-node
    length budget; the depot is always feasible."""
    cur_loc = gather_by_index(td["locs"], td["current_node"])[..., None, :]
    exceeds = td["tour_length"][..., None] + (td["locs"] - cur_loc).norm(p=2, dim=-1) > td["max_length"]
    mask = td["visited"] | td["visited"][..., 0:1] | exceeds
    action_mask = ~mask
    action_mask[..., 0] = 1
    return action_mask


def op_reset(depot, locs, prize, max_length) -> State:
    """op/env.py:110-152 (+ merge): per-node budget = max_length - dist(depot, node) - 1e-6."""
    B, n, _ = locs.shape
    lw = torch.cat((depot[:, None, :], locs), -2)
    td = {
        "locs": lw,
        "prize": F.pad(prize, (1, 0), mode="constant", value=0),
        "tour_length": torch.zeros(B),
        "max_length": max_length[..., None] - (depot[..., None, :] - lw).norm(p=2, dim=-1) - 1e-6,
        "current_node": torch.zeros(B, 1, dtype=torch.int64),
        "visited": torch.zeros(B, n + 1, dtype=torch.bool),
        "current_total_prize": torch.zeros(B),
        "i": torch.zeros(B, dtype=torch.int64),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }
    td["action_mask"] = op_action_mask(td)
    return td


def op_step(td: State) -> State:
    """op/env.py:74-108.  done = back at the depot after the first step, shape [B]."""
    cur = td["action"][:, None]
    prev_loc = gather_by_index(td["locs"], td["current_node"])
    cur_loc = gather_by_index(td["locs"], cur)
    length = td["tour_length"] + (cur_loc - prev_loc).norm(p=2, dim=-1)
    total = td["current_total_prize"] + gather_by_index(td["prize"], cur, dim=-1)
    visited = td["visited"].scatter(-1, cur, 1)
    done = (cur.squeeze(-1) == 0) & (td["i"] > 0)
    td.update({"tour_length": length, "current_node": cur, "visited": visited,
               "current_total_prize": total, "i": td["i"] + 1,
               "reward": torch.zeros_like(done), "done": done})
    td["action_mask"] = op_action_mask(td)
    return td


def op_check(locs_with_depot, max_length, actions) -> None:
    """op/env.py:184-215 (add_distance_to_depot=True)."""
    sa = actions.sort(1)[0]
    assert ((sa[:, 1:] == 0) | (sa[:, 1:] > sa[:, :-1])).all(), "Duplicates"
    length = tour_length(gather_by_index(locs_with_depot, actions))
    ml = max_length + (locs_with_depot[..., 0:1, :] - locs_with_depot).norm(p=2, dim=-1) + 1e-6
    assert (length[..., None] <= ml + 1e-5).all(), "Max length exceeded"


def op_reward(prize_with_depot, actions):
    """op/env.py:172-182."""
    if actions.size(-1) == 1:
        assert (actions == 0).all(), "If all length 1 tours, they should be zero"
        return torch.zeros(actions.size(0))
    return prize_with_depot.gather(1, actions).sum(-1)


# =====================================================================================
# PCTSP / SPCTSP  (envs/routing/pctsp/env.py, envs/routing/spctsp/env.py)
# =====================================================================================
def pctsp_action_mask(td: State):
    """pctsp/env.py:141-148: no node twice, nothing after the return to the depot; the depot is closed while
    the collected prize is below 1 and customers remain."""
    mask = td["visited"] | td["visited"][..., 0:1]
    mask[..., 0] = (td["cur_total_prize"] < 1.0) & (
        td["visited"][..., 1:].int().sum(-1) < td["visited"][..., 1:].size(-1))
    return ~(mask > 0)


def pctsp_reset(depot, locs, penalty, deterministic_prize, stochastic_prize, prize_required: float = 1.0,
                stochastic: bool = False) -> State:
    """pctsp/env.py:103-139 (+ merge).  spctsp: the collected (real) prize is the stochastic one, the encoder
    still sees the expected prize (spctsp/env.py:8-17)."""
    B, n, _ = locs.shape
    real = stochastic_prize if stochastic else deterministic_prize
    td = {
        "locs": torch.cat([depot[..., None, :], locs], dim=-2),
        "current_node": torch.zeros(B, dtype=torch.int64),
        "expected_prize": deterministic_prize,
        "real_prize": torch.cat([torch.zeros_like(real[..., :1]), real], dim=-1),
        "penalty": F.pad(penalty, (1, 0), mode="constant", value=0),
        "cur_total_prize": torch.zeros(B),
        "cur_total_penalty": penalty.sum(-1),
        "visited": torch.zeros(B, n + 1, dtype=torch.bool),
        "prize_required": torch.full((B,), prize_required),
        "i": torch.zeros(B, dtype=torch.int64),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }
    td["action_mask"] = pctsp_action_mask(td)
    return td


def pctsp_step(td: State) -> State:
    """pctsp/env.py:69-101."""
    cur = td["action"]
    prize = td["cur_total_prize"] + gather_by_index(td["real_prize"], cur)
    pen = td["cur_total_penalty"] + gather_by_index(td["penalty"], cur)
    visited = td["visited"].scatter(-1, cur[..., None], 1)
    done = (td["i"] > 0) & (cur == 0)
    td.update({"current_node": cur, "cur_total_prize": prize, "cur_total_penalty": pen, "visited": visited,
               "i": td["i"] + 1, "reward": torch.zeros_like(done), "done": done})
    td["action_mask"] = pctsp_action_mask(td)
    return td


def pctsp_check(td: State, actions) -> None:
    """pctsp/env.py:176-200."""
    sa = actions.sort(1)[0]
    assert ((sa[..., 1:] == 0) | (sa[..., 1:] > sa[..., :-1])).all(), "Duplicates"
    p = td["real_prize"].gather(1, actions)
    assert ((p.sum(-1) >= 1 - 1e-5) | (sa.size(-1) - (sa == 0).int().sum(-1) == (td["locs"].size(-2) - 1))
            ).all(), "Total prize does not satisfy min total prize"


def pctsp_reward(td: State, actions):
    """pctsp/env.py:150-174: saved penalties - (tour length from and to the depot + all penalties)."""
    if actions.size(-1) == 1:
        assert (actions == 0).all(), "If all length 1 tours, they should be zero"
        return torch.zeros(actions.size(0))
    ordered = torch.cat([td["locs"][..., 0:1, :], gather_by_index(td["locs"], actions)], dim=1)
    saved = td["penalty"].gather(1, actions)
    return saved.sum(-1) - (tour_length(ordered) + td["penalty"][..., 1:].sum(-1))


# =====================================================================================
# ATSP  (envs/routing/atsp/env.py) - env only: the AM init embedding registered for "atsp" needs td["locs"]
# =====================================================================================
def atsp_reset(cost_matrix) -> State:
    """atsp/env.py:92-112 (+ merge)."""
    B, N, _ = cost_matrix.shape
    cur = torch.zeros(B, 1, dtype=torch.int64)
    return {
        "cost_matrix": cost_matrix,
        "first_node": cur,
        "current_node": cur,
        "i": torch.zeros(B, 1, dtype=torch.int64),
        "action_mask": torch.ones(B, N, dtype=torch.bool),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }


def atsp_step(td: State) -> State:
    """atsp/env.py:64-90: first-step test on element 0 of i (utils/ops.py batch_to_scalar), done [B]."""
    cur = td["action"]
    first = cur if td["i"].reshape(-1)[0].item() == 0 else td["first_node"]
    avail = td["action_mask"].scatter(-1, cur.unsqueeze(-1).expand_as(td["action_mask"]), 0)
    done = torch.count_nonzero(avail, dim=-1) <= 0
    td.update({"first_node": first, "current_node": cur, "i": td["i"] + 1, "action_mask": avail,
               "reward": torch.zeros_like(done), "done": done})
    return td


def atsp_check(actions) -> None:
    """atsp/env.py:164-171."""
    ok = (torch.arange(actions.size(1)).view(1, -1).expand_as(actions) == actions.sort(1)[0]).all()
    assert ok, "Invalid tour"


def atsp_reward(cost_matrix, actions):
    """atsp/env.py:152-162: minus the sum of the cost-matrix entries along the closed tour."""
    tgt = torch.roll(actions, -1, dims=1)
    bidx = torch.arange(cost_matrix.shape[0]).unsqueeze(1)
    return -cost_matrix[bidx, actions, tgt].sum(-1)


# =====================================================================================
# SDVRP  (envs/routing/sdvrp/env.py; dynamic embedding nn/env_embeddings/dynamic.py:61-79)
# =====================================================================================
def sdvrp_action_mask(td: State):
    """sdvrp/env.py:132-140."""
    mask_loc = (td["demand_with_depot"][..., 1:] == 0) | (td["used_capacity"] >= td["vehicle_capacity"])
    mask_depot = (td["current_node"] == 0).squeeze(-1) & ((mask_loc == 0).int().sum(-1) > 0)
    return ~torch.cat((mask_depot[..., None], mask_loc), -1)


def sdvrp_reset(depot, locs, demand, vehicle_capacity: float = 1.0) -> State:
    """sdvrp/env.py:103-130 (+ merge)."""
    B = locs.shape[0]
    td = {
        "locs": torch.cat((depot[..., None, :], locs), -2),
        "demand": demand,
        "demand_with_depot": torch.cat((torch.zeros_like(demand[..., 0:1]), demand), -1),
        "current_node": torch.zeros(B, 1, dtype=torch.int64),
        "used_capacity": torch.zeros(B, 1),
        "vehicle_capacity": torch.full((B, 1), vehicle_capacity),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    }
    td["action_mask"] = sdvrp_action_mask(td)
    return td


def sdvrp_step(td: State) -> State:
    """sdvrp/env.py:63-101: deliver min(remaining demand, remaining capacity), nodes may be revisited."""
    cur = td["action"][:, None]
    sel = gather_by_index(td["demand_with_depot"], cur, dim=-1, squeeze=False)[..., :1]
    delivered = torch.min(sel, td["vehicle_capacity"] - td["used_capacity"])
    used = (td["used_capacity"] + delivered) * (cur != 0).float()
    dwd = td["demand_with_depot"].scatter_add(-1, cur, -delivered)
    done = ~(dwd > 0).any(-1)
    td.update({"demand_with_depot": dwd, "current_node": cur, "used_capacity": used,
               "reward": torch.zeros_like(done), "done": done})
    td["action_mask"] = sdvrp_action_mask(td)
    return td


def sdvrp_check(demand, vehicle_capacity, actions) -> None:
    """sdvrp/env.py:142-164."""
    B = demand.size(0)
    demands = torch.cat((-vehicle_capacity, demand), 1)
    rng = torch.arange(B)
    used = torch.zeros_like(demand[..., 0])
    a_prev = None
    for a in actions.transpose(0, 1):
        assert a_prev is None or (demands[((a_prev == 0) & (a == 0)), :] == 0).all(), \
            "Cannot visit depot twice if any nonzero demand"
        d = torch.min(demands[rng, a], vehicle_capacity.squeeze(-1) - used)
        demands[rng, a] -= d
        used += d
        used[a == 0] = 0
        a_prev = a
    assert (demands == 0).all(), "All demand must be satisfied"


# =====================================================================================
# SLAP  (envs/warehousing/slap/env.py)
# =====================================================================================
def slap_reset(locs, assignment, n_products: int, extra: Optional[State] = None) -> State:
    """slap/env.py:95-129 (+ merge: generator keys survive in the td)."""
    B, N, _ = locs.shape
    mask = torch.ones(B, N, dtype=torch.bool)
    mask[..., 0] = False
    td = dict(extra or {})
    td.update({
        "locs": locs,
        "assignment": assignment,
        "to_choose": torch.arange(n_products, dtype=torch.float32).unsqueeze(0).repeat(B, 1),
        "i": torch.zeros(B, 1, dtype=torch.int64),
        "ratio": torch.zeros(B, N),
        "action_mask": mask,
        "reward": torch.zeros(B, 1, dtype=torch.float32),
        "done": torch.zeros(B, 1, dtype=torch.bool),
        "terminated": torch.zeros(B, 1, dtype=torch.bool),
    })
    td["_n_products"] = n_products
    return td


def slap_step(td: State) -> State:
    """slap/env.py:38-93.  The reference clears one mask bit per instance in a python loop
    (:61-62); the scatter below is the same update."""
    product = td["to_choose"][..., 0].to(torch.int)
    to_choose = td["to_choose"][..., 1:]
    loc = td["action"].to(torch.int)
    B = loc.shape[0]
    assignment = td["assignment"].clone()
    assignment[torch.arange(B), product.long()] = loc
    done = td["i"] == td["_n_products"] - 1
    mask = td["action_mask"].clone()
    mask[torch.arange(B), loc.long()] = False
    td.update({"assignment": assignment, "to_choose": to_choose, "action_mask": mask,
               "i": td["i"] + 1, "reward": torch.zeros_like(done), "done": done})
    return td


def slap_reward(locs, assignment, picklist):
    """slap/env.py:131-143: sum over orders of minus the closed Euclidean tour through the
    locations assigned to the order's SKUs, accumulated order by order in fp32."""
    B = picklist.shape[0]
    total = torch.full((B,), 0, dtype=torch.float32)
    rows = torch.arange(B).unsqueeze(1)
    for o in range(picklist.shape[1]):
        subset = assignment[rows, picklist[:, o, :]].to(torch.int)
        total += -tour_length(locs[rows, subset.long()])
    return total


# =====================================================================================
# AM decoder  (models/zoo/am/decoder.py, models/nn/attention.py, utils/decoding.py)
# =====================================================================================
# weight dict keys (names follow policy.state_dict()):
#   "Wnode" [3E,E]  decoder.project_node_embeddings.weight
#   "Wfixed" [E,E]  decoder.project_fixed_context.weight
#   "Wctx"  [E,2E] (tsp) / [E,E+1] (cvrp) / absent (slap)  decoder.context_embedding.project_context.weight
#   "Wph"   [2E]    decoder.context_embedding.W_placeholder (tsp only)
#   "Wout"  [E,E]   decoder.pointer.project_out.weight
NUM_HEADS = 8


def precompute_cache(h, W, use_graph_context: bool = True):
    """am/decoder.py:208-237."""
    gk, gv, lk = F.linear(h, W["Wnode"]).chunk(3, dim=-1)
    gctx = F.linear(h.mean(1), W["Wfixed"]) if use_graph_context else 0
    return {"h": h, "gctx": gctx, "gk": gk, "gv": gv, "lk": lk}


def context_embedding(env_name: str, cache, td: State, W):
    """tsp: upstream TSPContext, nn/env_embeddings/context.py:102-137 (step 0 is decided
    by element 0 of td['i'], :121).  cvrp: EnvContext.forward + VRPContext, :58-71,317-331.
    slap: examples/slap.py:50-63 (zeros)."""
    h = cache["h"]
    B = h.size(0)
    if env_name == "tsp":
        if td["i"][(0,) * td["i"].dim()].item() < 1:
            if td["first_node"].dim() == 1:
                x = W["Wph"][None, :].expand(B, -1)
            else:
                x = W["Wph"][None, None, :].expand(B, td["first_node"].size(1), -1)
        else:
            node_dim = (-1,) if td["first_node"].dim() == 1 else (td["first_node"].size(-1), -1)
            x = gather_by_index(
                h, torch.stack([td["first_node"], td["current_node"]], -1).view(B, -1)
            ).view(B, *node_dim)
        return F.linear(x, W["Wctx"])
    if env_name in ("cvrp", "sdvrp"):
        cur = gather_by_index(h, td["current_node"])
        state = (td["vehicle_capacity"] - td["used_capacity"]).to(cur.dtype)   # (no-op in fp32; fp64 decoder checks)
        return F.linear(torch.cat([cur, state], -1), W["Wctx"])
    if env_name == "op":        # OPContext, nn/env_embeddings/context.py:385-397
        cur = gather_by_index(h, td["current_node"])
        state = (td["max_length"][..., 0] - td["tour_length"])[..., None].to(cur.dtype)
        return F.linear(torch.cat([cur, state], -1), W["Wctx"])
    if env_name in ("pctsp", "spctsp"):     # PCTSPContext, context.py:368-382
        cur = gather_by_index(h, td["current_node"])
        state = torch.clamp(td["prize_required"] - td["cur_total_prize"], min=0)[..., None].to(cur.dtype)
        return F.linear(torch.cat([cur, state], -1), W["Wctx"])
    if env_name == "slap":
        return h.new_zeros(B, h.size(-1))
    raise ValueError(env_name)


def pointer_logits(q, cache, mask, W):
    """nn/attention.py:263-309: masked 8-head SDPA, project_out, logits / sqrt(E)."""
    def heads(x):  # "... g (h s) -> ... h g s"
        return x.view(*x.shape[:-1], NUM_HEADS, -1).transpose(-3, -2)

    qh, kh, vh = heads(q), heads(cache["gk"]), heads(cache["gv"])
    am = mask.unsqueeze(1) if mask.ndim == 3 else mask.unsqueeze(1).unsqueeze(2)
    o = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=am)
    o = o.transpose(-3, -2).reshape(*q.shape)  # "... h n g -> ... n (h g)"
    glimpse = F.linear(o, W["Wout"])
    logits = torch.bmm(glimpse, cache["lk"].transpose(-2, -1)).squeeze(-2) / math.sqrt(glimpse.size(-1))
    assert not torch.isnan(logits).any(), "Logits contain NaNs"
    return logits


def decoder_forward(env_name, td_flat: State, cache, W, num_starts: int):
    """am/decoder.py:162-200: returns (logits, mask), flattened start-major for multistart."""
    td = td_flat
    if env_name == "sdvrp" and num_starts > 1:
        # has_dyn_emb_multi_start, am/decoder.py:176-187: keys and values are not shared across starts (the episodes modify
        # them), so the CACHE is batchified to [S*B, ...] and the state stays flat - S*B independent single-start rows
        cache = {k: (batchify(v, num_starts) if isinstance(v, torch.Tensor) else v) for k, v in cache.items()}
        num_starts = 0
    if num_starts > 1:
        td = {k: unbatchify(v, num_starts) for k, v in td_flat.items() if isinstance(v, torch.Tensor)}
    gctx = cache["gctx"]
    if num_starts > 1 and isinstance(gctx, torch.Tensor):
        gctx = gctx.unsqueeze(1)
    q = context_embedding(env_name, cache, td, W) + gctx
    if q.ndim == 2:
        q = q.unsqueeze(1)
    mask = td["action_mask"]
    if env_name == "sdvrp":     # SDVRPDynamicEmbedding, dynamic.py:61-79 + _compute_kvl, am/decoder.py:148-160
        d = td["demand_with_depot"][..., None].clone().to(cache["gk"].dtype)
        d[..., 0, :] = 0
        k_dyn, v_dyn, l_dyn = F.linear(d, W["Wdyn"]).chunk(3, dim=-1)
        cache = dict(cache, gk=cache["gk"] + k_dyn, gv=cache["gv"] + v_dyn, lk=cache["lk"] + l_dyn)
    logits = pointer_logits(q, cache, mask, W)
    if num_starts > 1:
        S = num_starts
        logits = logits.permute(1, 0, 2).reshape(-1, logits.size(-1))  # "b s l -> (s b) l"
        mask = mask.permute(1, 0, 2).reshape(-1, mask.size(-1))
    return logits, mask


def process_logits(logits, mask, temperature: float = 1.0, tanh_clipping: float = 10.0):
    """utils/decoding.py:141-191 (top-k / top-p are out of scope)."""
    if tanh_clipping > 0:
        logits = torch.tanh(logits) * tanh_clipping
    logits = logits.masked_fill(~mask, float("-inf"))
    logits = logits / temperature
    return F.log_softmax(logits, dim=-1)


def select_action(logprobs, mask, decode_type: str, given=None):
    """utils/decoding.py:371-397, 410-499."""
    if decode_type == "evaluate":
        return given
    if "greedy" in decode_type:
        sel = logprobs.argmax(dim=-1)
    else:
        probs = logprobs.exp()
        sel = torch.multinomial(probs, 1).squeeze(1)
        while (~mask).gather(1, sel.unsqueeze(-1)).any():
            sel = probs.multinomial(1).squeeze(1)
    assert not (~mask).gather(1, sel.unsqueeze(-1)).any(), "infeasible action selected"
    return sel


_STEP = {"tsp": tsp_step, "cvrp": cvrp_step, "slap": slap_step, "op": op_step, "pctsp": pctsp_step,
         "spctsp": pctsp_step, "atsp": atsp_step, "sdvrp": sdvrp_step}


def am_rollout(env_name: str, td: State, h, W, decode_type: str = "greedy",
               num_starts: Optional[int] = None, actions=None, temperature: float = 1.0,
               tanh_clipping: float = 10.0, use_graph_context: bool = True,
               num_loc: Optional[int] = None, return_all_logp: bool = False,
               check_solution: bool = True, max_steps: int = 1_000_000, start_nodes=None):
    """models/common/constructive/base.py:196-276 with the decoding hooks of
    utils/decoding.py:265-369.  ``td`` is a reset state (tsp_reset / cvrp_reset /
    slap_reset); ``h`` the encoder output [B,N,E]."""
    td = dict(td)
    if actions is not None:
        decode_type = "evaluate"
    multistart = "multistart" in decode_type
    logps, acts = [], []
    S = 0
    if multistart:  # decoding.py:265-313
        S = num_starts if num_starts is not None else num_starts_for(env_name, td["action_mask"])
        B0 = td["action_mask"].shape[0]
        nl = num_loc if num_loc is not None else (td["locs"].shape[1] - (0 if env_name == "tsp" else 1))
        # start_nodes [S*B]: what select_start_nodes_fn returned (utils/decoding.py:287-291), e.g. the op branch of
        # utils/ops.py:166-176 that resamples the starts from the feasible nodes with torch.multinomial
        a0 = select_start_nodes(env_name, B0, S, nl) if start_nodes is None else torch.as_tensor(start_nodes, dtype=torch.int64)
        td = {k: (batchify(v, S) if isinstance(v, torch.Tensor) else v) for k, v in td.items()}
        td["action"] = a0
        td = _STEP[env_name](td)
        logps.append(torch.zeros_like(td["action_mask"]) if return_all_logp else torch.zeros_like(a0))
        acts.append(a0)
    cache = precompute_cache(h, W, use_graph_context)
    step = 0
    while not td["done"].all():
        logits, mask = decoder_forward(env_name, td, cache, W, S)
        lp = process_logits(logits, mask, temperature, tanh_clipping)
        given = actions[..., step] if actions is not None else None
        sel = select_action(lp, mask, decode_type, given)
        logps.append(lp if return_all_logp else gather_by_index(lp, sel, dim=1))
        acts.append(sel)
        td["action"] = sel
        td = _STEP[env_name](td)
        step += 1
        if step > max_steps:
            break
    logprobs = torch.stack(logps, 1)
    out_actions = torch.stack(acts, 1)
    if env_name == "tsp":
        reward = tsp_reward(td["locs"], out_actions, check_solution)
    elif env_name == "cvrp":
        if check_solution:
            cvrp_check(td["demand"], td["vehicle_capacity"], out_actions)
        reward = cvrp_reward(td["locs"], out_actions)
    elif env_name == "sdvrp":
        if check_solution:
            sdvrp_check(td["demand"], td["vehicle_capacity"], out_actions)
        reward = cvrp_reward(td["locs"], out_actions)
    elif env_name == "op":
        if check_solution:
            op_check(td["locs"], td["max_length"], out_actions)
        reward = op_reward(td["prize"], out_actions)
    elif env_name in ("pctsp", "spctsp"):
        if check_solution:
            pctsp_check(td, out_actions)
        reward = pctsp_reward(td, out_actions)
    else:
        reward = slap_reward(td["locs"], td["assignment"], td["picklist"])
    lp_sel = logprobs.gather(-1, out_actions.unsqueeze(-1)).squeeze(-1) if logprobs.dim() == 3 else logprobs
    assert (lp_sel > -1000).all(), "Logprobs should not be -inf, check sampling procedure!"
    return {"reward": reward, "log_likelihood": lp_sel.sum(1), "logprobs": logprobs,
            "actions": out_actions, "td": td}
