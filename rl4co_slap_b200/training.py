"""Training step for the rollout path: the rollout kernels produce the
trajectories (inference only), this module re-evaluates their log-likelihood DIFFERENTIABLY so that
REINFORCE can back-propagate into the decoder weights and, through ``h``, into the encoder.

Because the actions are known, the autoregressive loop of ``ConstructivePolicy.forward``
(rl4co/models/common/constructive/base.py:196-276, ``decode_type="evaluate"``, :206-207) collapses
into ONE batched decoder pass over all steps: every step's context query, mask, glimpse, pointer
logits and log-softmax are computed together with plain PyTorch ops (autograd does the backward),
in the op order of rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309 and
rl4co/utils/decoding.py:141-191.  The environment quantities that feed the decoder (action masks,
remaining capacity) are replayed from the actions with the reference's own arithmetic
(rl4co/envs/routing/tsp/env.py:67-93, cvrp/env.py:73-149, rl4co/envs/warehousing/slap/env.py:38-93).

``reinforce_step`` is ``REINFORCE.shared_step`` + ``calculate_loss``
(rl4co/models/rl/reinforce/reinforce.py:62-115): sample with the kernel, baseline, loss
``((reward - bl) * ll).mean()`` with the sign convention of the reference, backward, optimiser step.
Pure torch: runs on CUDA for training and on CPU for the parity tests against the oracle.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F


def _exclusive_visited(actions: torch.Tensor, N: int) -> torch.Tensor:
    """[R,T] int64 -> bool [R,T,N]: node n was chosen at some step < t."""
    onehot = F.one_hot(actions, N).to(torch.int32)
    return (onehot.cumsum(1) - onehot) > 0


def replay_masks(env_name: str, td, actions: torch.Tensor, S: int = 1):
    """Action masks [R,T,N] (True = feasible), per-step validity [R,T] (False once the episode is done) and, for
    cvrp / op, the used capacity / tour length [R,T] seen by the context of every step, replayed from ``actions`` [R,T]
    (R = S * B, start-major as utils/ops.py:11-34)."""
    R, T = actions.shape
    N = td["locs"].shape[1]
    dev = actions.device
    if env_name == "tsp":       # tsp/env.py:78-81: the mask is "not visited yet"
        return ~_exclusive_visited(actions, N), torch.ones(R, T, dtype=torch.bool, device=dev), None
    if env_name == "slap":      # slap/env.py:95-129,86-88: depot never assignable, a location once
        mask = ~_exclusive_visited(actions, N)
        mask[:, :, 0] = False
        return mask, torch.ones(R, T, dtype=torch.bool, device=dev), None
    if env_name not in ("cvrp", "op", "pctsp", "spctsp"):
        raise NotImplementedError(f"{env_name}: the differentiable teacher-forced pass covers tsp / cvrp / slap / op / pctsp / "
                                  "spctsp (sdvrp's keys and values change every step with the dynamic embedding)")
    if env_name == "op":
        return _replay_masks_op(td, actions, S)
    if env_name in ("pctsp", "spctsp"):
        return _replay_masks_pctsp(td, actions, S)
    # cvrp/env.py:73-149 - sequential in the used capacity (fp32, the reference's order of operations)
    demand = td["demand"].float()
    cap = td["vehicle_capacity"].float().reshape(-1, 1)
    if S > 1:
        demand, cap = demand.repeat(S, 1), cap.repeat(S, 1)
    used = torch.zeros(R, 1, device=dev)
    visited = torch.zeros(R, N, dtype=torch.bool, device=dev)
    cur = torch.zeros(R, dtype=torch.int64, device=dev)
    masks, useds, valid = [], [], []
    done = torch.zeros(R, dtype=torch.bool, device=dev)
    rows = torch.arange(R, device=dev)
    for t in range(T):
        exceeds = demand + used > cap                                   # :139
        mask_loc = visited[:, 1:] | exceeds                            # :142
        mask_depot = (cur == 0) & ((~mask_loc).sum(-1) > 0)            # :145
        masks.append(~torch.cat((mask_depot[:, None], mask_loc), -1))  # :148
        useds.append(used[:, 0])
        valid.append(~done)
        a = actions[:, t]
        sel = demand.gather(1, (a - 1).clamp(0, N - 2)[:, None])       # :76-79
        used = (used + sel) * (a != 0).float()[:, None]                # :83-85
        visited = visited.clone()
        visited[rows, a] = True                                        # :88
        done = done | (visited.sum(-1) == N)                           # :91
        cur = a
    return torch.stack(masks, 1), torch.stack(valid, 1), torch.stack(useds, 1)


def _leg(d: torch.Tensor) -> torch.Tensor:
    """[..., 2] -> [...] Euclidean length with the rounding of torch's CPU vector norm on either device
    (sqrt(fma(dy, dy, dx dx)); on CUDA addcmul is the fused multiply-add)."""
    if not d.is_cuda:
        return d.norm(p=2, dim=-1)
    return torch.sqrt(torch.addcmul(d[..., 0] * d[..., 0], d[..., 1], d[..., 1]))


def _replay_masks_op(td, actions: torch.Tensor, S: int):
    """op/env.py:74-170 replayed from the actions: masks, validity and the tour length seen by every step."""
    R, T = actions.shape
    locs, budget = td["locs"].float(), td["max_length"].float()
    if S > 1:
        locs, budget = locs.repeat(S, 1, 1), budget.repeat(S, 1)
    N, dev = locs.shape[1], actions.device
    rows = torch.arange(R, device=dev)
    length = torch.zeros(R, device=dev)
    visited = torch.zeros(R, N, dtype=torch.bool, device=dev)
    cur = torch.zeros(R, dtype=torch.int64, device=dev)
    done = torch.zeros(R, dtype=torch.bool, device=dev)
    masks, lengths, valid = [], [], []
    for t in range(T):
        exceeds = length[:, None] + _leg(locs - locs[rows, cur][:, None, :]) > budget     # :159-162
        m = ~(visited | visited[:, 0:1] | exceeds)                                        # :163-165
        m[:, 0] = True                                                                    # :168
        masks.append(m)
        lengths.append(length)
        valid.append(~done)
        a = actions[:, t]
        length = length + _leg(locs[rows, a] - locs[rows, cur])                           # :78-81
        visited = visited.clone()
        visited[rows, a] = True
        done = done | ((a == 0) & (t > 0))                                                # :92
        cur = a
    return torch.stack(masks, 1), torch.stack(valid, 1), torch.stack(lengths, 1)


def _replay_masks_pctsp(td, actions: torch.Tensor, S: int):
    """pctsp/env.py:69-148 replayed from the actions: masks, validity and the collected prize seen by every step."""
    R, T = actions.shape
    prize = td["real_prize"].float()
    if S > 1:
        prize = prize.repeat(S, 1)
    N, dev = prize.shape[1], actions.device
    rows = torch.arange(R, device=dev)
    total = torch.zeros(R, device=dev)
    visited = torch.zeros(R, N, dtype=torch.bool, device=dev)
    done = torch.zeros(R, dtype=torch.bool, device=dev)
    masks, totals, valid = [], [], []
    for t in range(T):
        m = visited | visited[:, 0:1]                                                     # :143
        m[:, 0] = (total < 1.0) & (visited[:, 1:].int().sum(-1) < N - 1)                  # :144-146
        masks.append(~m)
        totals.append(total)
        valid.append(~done)
        a = actions[:, t]
        total = total + prize[rows, a]                                                    # :73-75
        visited = visited.clone()
        visited[rows, a] = True
        done = done | ((a == 0) & (t > 0))                                                # :84
    return torch.stack(masks, 1), torch.stack(valid, 1), torch.stack(totals, 1)


def teacher_forced_log_likelihood(policy, td, env_name: str, h: torch.Tensor, actions: torch.Tensor,
                                  num_starts: int = 0, temperature: Optional[float] = None,
                                  tanh_clipping: Optional[float] = None, mask_logits: Optional[bool] = None,
                                  return_sum: bool = True) -> torch.Tensor:
    """Differentiable log-likelihood of ``actions`` [R,T] under the policy's decoder given the encoder
    output ``h`` [B,N,E] (R = B, or S*B start-major with ``num_starts`` = S POMO starts whose forced first
    action carries log-prob 0, utils/decoding.py:287-305).  Returns [R] (or [R,T])."""
    dec = policy.decoder
    E, H = dec.embed_dim, dec.num_heads
    B, N, _ = h.shape
    S = max(int(num_starts), 1)
    R, T = actions.shape
    assert R == B * S, f"actions has {R} rows, expected {B} x {S}"
    temperature = float(policy.temperature if temperature is None else temperature)
    tanh_clipping = float(policy.tanh_clipping if tanh_clipping is None else tanh_clipping)
    mask_logits = bool(policy.mask_logits if mask_logits is None else mask_logits)
    multistart = num_starts > 0
    dev = h.device

    mask, valid, used = replay_masks(env_name, td, actions, S)           # [R,T,N], [R,T], [R,T]
    # ---- cache (am/decoder.py:208-237)
    gk, gv, lk = dec.project_node_embeddings(h).chunk(3, dim=-1)          # [B,N,E] each
    gctx = dec.project_fixed_context(h.mean(1)) if dec.use_graph_context else None   # [B,E]

    # rollouts of instance b as extra queries: [R,T] start-major -> [B, L = S*T]
    def per_instance(x):                                                  # [R,T,...] -> [B,S*T,...]
        return x.reshape(S, B, T, *x.shape[2:]).transpose(0, 1).reshape(B, S * T, *x.shape[2:])
    a_bl = per_instance(actions)                                          # [B,L]
    L = S * T
    prev = torch.cat((torch.zeros(R, 1, dtype=torch.int64, device=dev), actions[:, :-1]), 1)   # current node of step t
    prev_bl = per_instance(prev)

    # ---- context embedding (env_embeddings/context.py:47-71,102-137,317-331; examples/slap.py:50-63)
    Wc = dec.context_embedding.project_context.weight
    gather = lambda src, idx: src.gather(1, idx[..., None].expand(-1, -1, E))
    if env_name == "tsp":
        first_bl = per_instance(actions[:, :1].expand(-1, T))
        ctx = gather(F.linear(h, Wc[:, :E]), first_bl) + gather(F.linear(h, Wc[:, E:]), prev_bl)
        if not multistart:      # step 0: no first / current node yet -> W_placeholder (context.py:111-121)
            ctx0 = F.linear(dec.context_embedding.W_placeholder, Wc)
            step0 = (torch.arange(L, device=dev) % T == 0)[None, :, None]
            ctx = torch.where(step0, ctx0.expand(B, L, E), ctx)
    elif env_name == "cvrp":
        cap = td["vehicle_capacity"].float().reshape(B, 1)
        rem = cap.repeat(S, 1) - used                                     # [R,T]  (context.py:326-329)
        ctx = gather(F.linear(h, Wc[:, :E]), prev_bl) + per_instance(rem)[..., None] * Wc[:, E]
    elif env_name == "op":            # [h_cur ; max_length[..., 0] - tour_length] (context.py:385-397)
        rem = td["max_length"].float()[:, :1].repeat(S, 1) - used
        ctx = gather(F.linear(h, Wc[:, :E]), prev_bl) + per_instance(rem)[..., None] * Wc[:, E]
    elif env_name in ("pctsp", "spctsp"):    # [h_cur ; max(prize_required - cur_total_prize, 0)] (context.py:368-382)
        rem = torch.clamp(td["prize_required"].float().reshape(B, 1).repeat(S, 1) - used, min=0)
        ctx = gather(F.linear(h, Wc[:, :E]), prev_bl) + per_instance(rem)[..., None] * Wc[:, E]
    else:
        ctx = torch.zeros(B, L, E, device=dev, dtype=h.dtype)
    q = ctx if gctx is None else ctx + gctx[:, None, :]                   # decoder.py:141-142

    # ---- pointer attention (nn/attention.py:263-309): masked multi-head glimpse, project_out, logits
    m_bl = per_instance(mask)                                             # [B,L,N] True = feasible
    heads = lambda x: x.reshape(B, -1, H, E // H).transpose(1, 2)         # [B,H,*,d]
    glimpse = F.scaled_dot_product_attention(heads(q), heads(gk), heads(gv), attn_mask=m_bl[:, None])
    glimpse = dec.pointer.project_out(glimpse.transpose(1, 2).reshape(B, L, E))
    logits = torch.matmul(glimpse, lk.transpose(1, 2)) / math.sqrt(E)     # [B,L,N]
    # ---- process_logits (utils/decoding.py:141-191)
    if tanh_clipping > 0:
        logits = torch.tanh(logits) * tanh_clipping
    if mask_logits:
        logits = logits.masked_fill(~m_bl, float("-inf"))
    logp = F.log_softmax(logits / temperature, dim=-1).gather(-1, a_bl[..., None])[..., 0]   # [B,L]
    keep = per_instance(valid)
    if multistart:              # the forced start node has log-prob 0
        keep = keep & (torch.arange(L, device=dev) % T != 0)[None]
    logp = torch.where(keep, logp, torch.zeros_like(logp))
    logp = logp.reshape(B, S, T).transpose(0, 1).reshape(R, T)            # back to start-major rows
    return logp.sum(1) if return_sum else logp


# ---- the same log-likelihood through the kernels, differentiable -------------------------------------------------
class _KernelLogLikelihood(torch.autograd.Function):
    """ll [R] of given tours from the rollout tables: forward = the rollout kernel in evaluate mode, backward =
    rl4co_am_evaluate_backward (csrc/am_backward.cu, recompute based).  Differentiable inputs: tables
    [n_tables, B*N, 128], gctx [B,128] or None, q0 [128] (tsp) or None, wcap [128] (cvrp) or None."""

    @staticmethod
    def _args(meta, tables, gctx, q0, wcap):
        from . import _lib
        a = _lib.AmRolloutArgs()
        a.env, a.mode = _lib.ENV_IDS[meta["env"]], _lib.MODE_EVALUATE
        a.B, a.S, a.multistart, a.num_loc = meta["B"], meta["S"], int(meta["S"] > 0), meta["num_loc"]
        a.N, a.T, a.n_tables = meta["N"], meta["T"], tables.shape[0]
        a.temperature, a.tanh_clipping, a.mask_logits = 1.0, meta["clip"], 1
        a.tables, a.gctx, a.q0, a.wcap = _lib.ptr(tables), _lib.ptr(gctx), _lib.ptr(q0), _lib.ptr(wcap)
        a.locs = _lib.ptr(meta["locs"])
        a.demand, a.vehicle_capacity = _lib.ptr(meta.get("demand")), _lib.ptr(meta.get("capacity"))
        a.actions_in = _lib.ptr(meta["actions"])
        return a

    @staticmethod
    def forward(ctx, tables, gctx, q0, wcap, meta):
        from . import _lib
        dev, R, T = tables.device, meta["actions"].shape[0], meta["T"]
        tables, gctx = tables.contiguous(), (gctx.contiguous() if gctx is not None else None)
        q0, wcap = (q0.contiguous() if q0 is not None else None), (wcap.contiguous() if wcap is not None else None)
        out = {"actions": torch.empty(R, T, dtype=torch.int64, device=dev), "ll": torch.empty(R, device=dev),
               "reward": torch.empty(R, device=dev), "status": _lib.new_status(dev)}
        a = _KernelLogLikelihood._args(meta, tables, gctx, q0, wcap)
        a.actions, a.ll, a.reward, a.status = _lib.ptr(out["actions"]), _lib.ptr(out["ll"]), _lib.ptr(out["reward"]), _lib.ptr(out["status"])
        _lib.check(_lib.load().rl4co_am_rollout(a, _lib.stream_ptr(dev)), "rl4co_am_rollout")
        _lib.raise_for_status(out["status"])
        ctx.meta = meta
        ctx.save_for_backward(tables, *(x for x in (gctx, q0, wcap) if x is not None))
        ctx.have = (gctx is not None, q0 is not None, wcap is not None)
        return out["ll"]

    @staticmethod
    def backward(ctx, grad_ll):
        from . import _lib
        saved = list(ctx.saved_tensors)
        tables = saved.pop(0)
        gctx, q0, wcap = (saved.pop(0) if have else None for have in ctx.have)
        dev = tables.device
        g_tables = torch.zeros_like(tables)
        g_gctx = torch.zeros_like(gctx) if gctx is not None else None
        g_q0 = torch.zeros_like(q0) if q0 is not None else None
        g_wcap = torch.zeros_like(wcap) if wcap is not None else None
        a = _KernelLogLikelihood._args(ctx.meta, tables, gctx, q0, wcap)
        _lib.check(_lib.load().rl4co_am_evaluate_backward(
            a, _lib.ptr(grad_ll.contiguous().float()), _lib.ptr(g_tables), _lib.ptr(g_gctx), _lib.ptr(g_q0), _lib.ptr(g_wcap),
            _lib.stream_ptr(dev)), "rl4co_am_evaluate_backward")
        return g_tables, g_gctx, g_q0, g_wcap, None


def kernel_backward_supported(policy, env_name: str, h: torch.Tensor) -> bool:
    from . import _lib
    return (h.is_cuda and env_name in ("tsp", "cvrp") and policy.tanh_clipping > 0 and policy.temperature == 1.0
            and bool(policy.mask_logits) and h.shape[1] <= int(_lib.load().rl4co_am_evaluate_backward_max_nodes()))


def kernel_log_likelihood(policy, td, env_name: str, h: torch.Tensor, actions: torch.Tensor, num_starts: int = 0) -> torch.Tensor:
    """The log-likelihood of :func:`teacher_forced_log_likelihood`, [R], with forward AND backward in the kernels.
    The table projection (am/decoder.py:208-237 with project_out and the context projection folded in, as in
    policy.AttentionModelDecoder._fold_weights) is built here with differentiable torch ops, so autograd carries the
    kernel's table gradients on into ``h`` (the encoder) and every decoder weight."""
    dec = policy.decoder
    E = dec.embed_dim
    B, N, _ = h.shape
    Wn = dec.project_node_embeddings.weight
    Wc = dec.context_embedding.project_context.weight
    parts = [Wn[:E], Wn[E:2 * E], dec.pointer.project_out.weight.t() @ Wn[2 * E:]]
    q0 = wcap = None
    if env_name == "tsp":
        parts += [Wc[:, E:], Wc[:, :E]]
        q0 = Wc @ dec.context_embedding.W_placeholder
    else:
        parts += [Wc[:, :E]]
        wcap = Wc[:, E]
    nt = len(parts)
    tables = F.linear(h.float(), torch.cat(parts, 0)).view(B * N, nt, E).permute(1, 0, 2).contiguous()
    gctx = dec.project_fixed_context(h.float().mean(1)) if dec.use_graph_context else None
    S = int(num_starts)
    meta = {"env": env_name, "B": B, "S": S, "N": N, "T": actions.shape[1], "clip": float(policy.tanh_clipping),
            "num_loc": N - (0 if env_name == "tsp" else 1), "locs": td["locs"].float().contiguous(),
            "actions": actions.contiguous()}
    if env_name == "cvrp":
        meta["demand"] = td["demand"].float().contiguous()
        meta["capacity"] = td["vehicle_capacity"].float().contiguous()
    return _KernelLogLikelihood.apply(tables, gctx, q0, wcap, meta)


def reinforce_step(policy, env, td, optimizer=None, baseline=None, decode_type: Optional[str] = None,
                   num_starts: Optional[int] = None, max_grad_norm: Optional[float] = 1.0,
                   backward: str = "auto") -> dict:
    """One REINFORCE training step (reinforce.py:62-115, pomo/model.py:87-111): roll out with the fused kernel
    (no autograd), re-evaluate the sampled trajectories' log-likelihood differentiably, loss
    ``((reward - baseline) * log_likelihood).mean()`` on the negated... exactly ``-(advantage * ll).mean()`` as in
    reinforce.py:103-106 with reward = -cost, backward, clip, optimiser step.  ``baseline``: None (no baseline),
    "shared" (POMO mean over the starts of an instance, baselines.py:57-61), a tensor [R], or a baseline object with the
    reference's ``eval(td, reward, env) -> (value, loss)`` (rl4co_slap_b200.baselines)."""
    decode_type = decode_type or policy.train_decode_type
    multistart = "multistart" in decode_type
    # ONE encoder pass over the whole batch with autograd (train-mode batch norm sees the full batch once and updates
    # its running statistics once); the rollout kernels sample from exactly this h, so the teacher-forced ll below is
    # the log-likelihood of the policy that drew the actions
    h, _ = policy.encoder(td)
    with torch.no_grad():
        out = policy(td, env, phase="train", decode_type=decode_type, num_starts=num_starts,
                     return_actions=True, hidden=h.detach())
    actions, reward = out["actions"], out["reward"]
    B = td["locs"].shape[0]
    S = actions.shape[0] // B if multistart else 0
    # differentiable re-evaluation of the sampled tours: the backward kernel where it applies (tsp / cvrp, standard logit
    # post-processing), the batched PyTorch pass otherwise; ``backward`` = "kernel" / "torch" forces one of them
    use_kernel = backward == "kernel" or (backward == "auto" and kernel_backward_supported(policy, env.name, h))
    if use_kernel:
        ll = kernel_log_likelihood(policy, td, env.name, h, actions, num_starts=S)
    else:
        ll = teacher_forced_log_likelihood(policy, td, env.name, h.float(), actions, num_starts=S)
    if baseline is None:
        bl = torch.zeros_like(reward)
    elif isinstance(baseline, str) and baseline == "shared":
        bl = reward.reshape(max(S, 1), B).mean(0, keepdim=True).expand(max(S, 1), B).reshape(-1)
    elif hasattr(baseline, "eval"):     # a baselines.py object (No / Shared / Exponential / Mean / Warmup / Rollout)
        if S > 1:   # POMO hands the baseline reward as [B,S] (pomo/model.py:99-107); back to the flat start-major order
            r2 = reward.view(S, B).t()
            bl, _ = baseline.eval(td, r2, env)
            bl = (torch.zeros_like(r2) + bl).t().reshape(-1)
        else:
            bl, _ = baseline.eval(td, reward, env)
    else:
        bl = baseline
    advantage = reward - bl
    loss = -(advantage * ll).mean()
    if optimizer is not None:
        optimizer.zero_grad(set_to_none=True)
        loss.backward()
        if max_grad_norm:
            torch.nn.utils.clip_grad_norm_(policy.parameters(), max_grad_norm)
        optimizer.step()
    return {"loss": loss.detach(), "reward": reward, "log_likelihood": ll.detach(), "actions": actions,
            "kernel_log_likelihood": out["log_likelihood"]}
