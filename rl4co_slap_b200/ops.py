"""Tensor helpers with the reference's names and semantics (rl4co/utils/ops.py:11-183).

These are layout / indexing glue around the kernels (views, expands, gathers); the heavy
lifting of the rollout path lives in csrc/.
"""
from __future__ import annotations

from typing import Union

import torch
from torch import Tensor

from .tensordict_compat import is_tensordict


def _batchify_single(x, repeats: int):
    """ops.py:11-16 - start-major repeat: out[s*B + b] = x[b]."""
    s = x.shape
    return x.expand(repeats, *s).contiguous().view(s[0] * repeats, *s[1:])


def batchify(x, shape: Union[tuple, int]):
    """ops.py:19-34."""
    shape = [shape] if isinstance(shape, int) else shape
    for s in reversed(shape):
        x = _batchify_single(x, s) if s > 0 else x
    return x


def _unbatchify_single(x, repeats: int):
    """ops.py:37-42 - [(S B), ...] -> [B, S, ...]."""
    s = x.shape
    return x.view(repeats, s[0] // repeats, *s[1:]).permute(1, 0, *range(2, len(s) + 1))


def unbatchify(x, shape: Union[tuple, int]):
    """ops.py:45-62."""
    shape = [shape] if isinstance(shape, int) else shape
    for s in reversed(shape):
        x = _unbatchify_single(x, s) if s > 0 else x
    return x


def gather_by_index(src: Tensor, idx: Tensor, dim: int = 1, squeeze: bool = True) -> Tensor:
    """ops.py:65-77."""
    expanded_shape = list(src.shape)
    expanded_shape[dim] = -1
    idx = idx.view(idx.shape + (1,) * (src.dim() - idx.dim())).expand(expanded_shape)
    squeeze = idx.size(dim) == 1 and squeeze
    return src.gather(dim, idx).squeeze(dim) if squeeze else src.gather(dim, idx)


def unbatchify_and_gather(x, idx: Tensor, n: int):
    """ops.py:80-85."""
    x = unbatchify(x, n)
    if is_tensordict(x):
        return x.gather(idx.dim(), idx) if hasattr(x, "gather") else x
    return gather_by_index(x, idx, dim=idx.dim())


def get_distance(x: Tensor, y: Tensor) -> Tensor:
    """ops.py:88-90."""
    return (x - y).norm(p=2, dim=-1)


def get_tour_length(ordered_locs: Tensor) -> Tensor:
    """ops.py:93-101."""
    return get_distance(torch.roll(ordered_locs, -1, dims=-2), ordered_locs).sum(-1)


def calculate_entropy(logprobs: Tensor) -> Tensor:
    """ops.py:114-122.  logprobs [batch, steps, actions]."""
    logprobs = torch.nan_to_num(logprobs, nan=0.0)
    entropy = -(logprobs.exp() * logprobs).sum(dim=-1).sum(dim=1)
    assert entropy.isfinite().all(), "Entropy is not finite"
    return entropy


def get_num_starts(td, env_name=None) -> int:
    """ops.py:126-136."""
    num_starts = td["action_mask"].shape[-1]
    if env_name == "pdp":
        num_starts = (num_starts - 1) // 2
    elif env_name in ["cvrp", "cvrptw", "sdvrp", "mtsp", "op", "pctsp", "spctsp"]:
        num_starts = num_starts - 1
    return num_starts


def select_start_nodes(td, env, num_starts: int) -> Tensor:
    """ops.py:139-163 (tsp / depot branches; the op / mtsp resampling branches are out of scope)."""
    num_loc = env.generator.num_loc if hasattr(env.generator, "num_loc") else 0xFFFFFFFF
    selected = torch.arange(num_starts, device=td.device).repeat_interleave(td.shape[0]) % num_loc
    if env.name in ["tsp", "atsp", "flp", "mcp"]:
        return selected
    return selected + 1


def get_best_actions(actions: Tensor, max_idxs: Tensor) -> Tensor:
    """ops.py:186-188."""
    actions = unbatchify(actions, max_idxs.shape[0])
    return actions.gather(0, max_idxs[..., None, None])
