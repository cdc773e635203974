"""Decoding utilities with the reference's names (rl4co/utils/decoding.py).

``decode_step`` is the standalone form of DecodingStrategy.step (process_logits + greedy /
sampling / evaluate + chosen log-prob) on caller-provided logits - one launch of
rl4co_decode_step.  ``rollout`` / ``random_policy`` are the env-only loop the reference's
tests use (decoding.py:81-109).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

DECODE_TYPES = {"greedy": _lib.MODE_GREEDY, "sampling": _lib.MODE_SAMPLING,
                "multistart_greedy": _lib.MODE_GREEDY, "multistart_sampling": _lib.MODE_SAMPLING,
                "evaluate": _lib.MODE_EVALUATE}


def decode_step(logits: torch.Tensor, mask: torch.Tensor, decode_type: str = "greedy",
                temperature: float = 1.0, tanh_clipping: float = 0.0, mask_logits: bool = True,
                action: Optional[torch.Tensor] = None, store_all_logp: bool = False,
                seed: Optional[int] = None, offset: int = 0):
    """utils/decoding.py:141-191,327-397 -> (logprobs [B] or [B,N], action [B]).
    Raises AssertionError("infeasible action selected") like decoding.py:377-379."""
    if decode_type not in DECODE_TYPES:
        raise NotImplementedError(f"decode_type {decode_type!r}")
    if not logits.is_cuda:
        raise RuntimeError("decode_step needs CUDA tensors (no CPU fallback)")
    mode = DECODE_TYPES[decode_type] if action is None else _lib.MODE_EVALUATE
    dev, (B, N) = logits.device, logits.shape
    out_a = torch.empty(B, dtype=torch.int64, device=dev)
    out_lp = torch.empty(B, dtype=torch.float32, device=dev)
    all_lp = torch.empty(B, N, dtype=torch.float32, device=dev) if store_all_logp else None
    status = _lib.new_status(dev)
    if seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if mode == _lib.MODE_SAMPLING else 0
    m = mask.view(torch.uint8) if mask.dtype == torch.bool else mask
    _lib.call("rl4co_decode_step", B, N, _lib.ptr(logits.contiguous(), torch.float32),
              _lib.ptr(m.contiguous(), torch.uint8), mode, float(temperature), float(tanh_clipping),
              int(mask_logits), _lib.ptr(action.contiguous(), torch.int64) if action is not None else None,
              seed, int(offset), _lib.ptr(out_a), _lib.ptr(out_lp), _lib.ptr(all_lp), _lib.ptr(status),
              _lib.stream_ptr(dev))
    if mode != _lib.MODE_EVALUATE:
        _lib.raise_for_status(status)
    return (all_lp if store_all_logp else out_lp), out_a


def get_log_likelihood(logprobs, actions=None, mask=None, return_sum: bool = True):
    """utils/decoding.py:39-65."""
    if actions is not None and logprobs.dim() == 3:
        logprobs = logprobs.gather(-1, actions.unsqueeze(-1)).squeeze(-1)
    if mask is not None:
        logprobs[~mask] = 0
    assert (logprobs > -1000).data.all(), "Logprobs should not be -inf, check sampling procedure!"
    return logprobs.sum(1) if return_sum else logprobs


def random_policy(td):
    """utils/decoding.py:81-85: uniformly random feasible action."""
    action = torch.multinomial(td["action_mask"].float(), 1).squeeze(-1)
    td.set("action", action)
    return td


def rollout(env, td, policy, max_steps: Optional[int] = None):
    """utils/decoding.py:88-109: env-only loop -> (reward, td, actions)."""
    max_steps = float("inf") if max_steps is None else max_steps
    actions, steps = [], 0
    while not td["done"].all():
        td = policy(td)
        actions.append(td["action"])
        td = env.step(td)["next"]
        steps += 1
        if steps > max_steps:
            break
    actions = torch.stack(actions, 1)
    return env.get_reward(td, actions), td, actions
