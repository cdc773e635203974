"""Rollout-order bookkeeping: how S starts and A augmented copies of B instances are laid out in one flat batch.

The kernels and the reference agree on ONE convention (rl4co/utils/ops.py:11-62): the repeat index is the slow
one.  Repeating a batch ``(A, S)`` times puts start ``s`` of copy ``a`` of instance ``b`` at row
``(s*A + a)*B + b``; ``unbatchify`` turns such a flat axis back into ``[B, A, S]``.  The public names are the
reference's (callers and tests use them); the implementations are written around that index formula: one tiling
and one view + axis reversal regardless of how many repeat levels there are.

Also here: the small tensor utilities of the same reference file that the host layer needs (index gather, closed
tour length, policy entropy) and the multistart conventions (number of starts, forced start nodes).
"""
from __future__ import annotations

import math
from typing import Sequence, Union

import torch
from torch import Tensor

from .tensordict_compat import is_tensordict

# environments whose node 0 is a depot that is never a POMO start (utils/ops.py:126-136, 158-163)
_DEPOT_ENVS = frozenset(("cvrp", "cvrptw", "sdvrp", "mtsp", "op", "pctsp", "spctsp"))
_NO_DEPOT_START_ENVS = frozenset(("tsp", "atsp", "flp", "mcp"))


def _levels(shape: Union[int, Sequence[int]]) -> list:
    """Repeat counts that actually repeat (the reference skips levels <= 0, utils/ops.py:31-33)."""
    levels = [shape] if isinstance(shape, int) else list(shape)
    return [int(n) for n in levels if n > 0]


def batchify(x, shape: Union[int, Sequence[int]]):
    """Tile ``x`` along its batch axis ``prod(shape)`` times, instance index fastest (utils/ops.py:11-34).
    Nested repeats compose into a single tiling, so ``(A, S)`` costs one copy, not two."""
    total = math.prod(_levels(shape))
    if total <= 1:
        return x
    lead = x.shape[0]
    tiled = x.expand(total, *x.shape).contiguous()          # [total, B, ...] (batch dims only for a TensorDict)
    return tiled.view(total * lead, *x.shape[1:])


def unbatchify(x, shape: Union[int, Sequence[int]]):
    """Inverse view of :func:`batchify`: ``[(S A B), ...] -> [B, A, S, ...]`` for ``shape = (A, S)``
    (utils/ops.py:37-62).  No copy: a view with the repeat axes reversed behind the instance axis."""
    levels = _levels(shape)
    if not levels:
        return x
    k = len(levels)
    lead = x.shape[0] // math.prod(levels)
    split = x.view(*levels[::-1], lead, *x.shape[1:])       # slowest repeat level first
    order = list(range(k, -1, -1)) + list(range(k + 1, k + 1 + len(x.shape) - 1))
    return split.permute(*order)


def gather_by_index(src: Tensor, idx: Tensor, dim: int = 1, squeeze: bool = True) -> Tensor:
    """``src`` indexed along ``dim`` by one index (or a list of them) per leading position
    (utils/ops.py:65-77): idx is broadcast over the trailing axes of ``src``."""
    index = idx.reshape(*idx.shape, *([1] * (src.dim() - idx.dim())))
    out = torch.take_along_dim(src, index, dim=dim)
    return out.squeeze(dim) if (squeeze and out.shape[dim] == 1) else out


def unbatchify_and_gather(x, idx: Tensor, n: int):
    """Pick entry ``idx[b]`` out of the ``n`` repeats of instance ``b`` (utils/ops.py:80-85)."""
    per_instance = unbatchify(x, n)
    if is_tensordict(per_instance):
        return per_instance.gather(idx.dim(), idx)
    return gather_by_index(per_instance, idx, dim=idx.dim())


def get_distance(x: Tensor, y: Tensor) -> Tensor:
    """Euclidean distance along the last axis (utils/ops.py:88-90)."""
    return torch.linalg.vector_norm(x - y, ord=2, dim=-1)


def get_tour_length(ordered_locs: Tensor) -> Tensor:
    """Length of the closed tour through ``ordered_locs [..., T, 2]`` (utils/ops.py:93-101)."""
    successor = ordered_locs.roll(shifts=-1, dims=-2)
    return get_distance(successor, ordered_locs).sum(dim=-1)


def calculate_entropy(logprobs: Tensor) -> Tensor:
    """Summed per-step entropy of the policy, ``logprobs [B, T, N]`` (utils/ops.py:114-122).  Masked actions carry
    -inf (and rows after ``done`` may carry NaN): both contribute zero, as in the reference, whose nan_to_num maps
    NaN to 0 and -inf to the lowest finite float, i.e. p * log p = 0 * finite."""
    p = logprobs.exp()
    terms = torch.where(torch.isfinite(logprobs) & (p > 0), p * logprobs, torch.zeros_like(p))
    entropy = -terms.sum(dim=(1, 2))
    assert torch.isfinite(entropy).all(), "Entropy is not finite"
    return entropy


def get_num_starts(td, env_name=None) -> int:
    """Default number of POMO starts = number of candidate start nodes (utils/ops.py:126-136)."""
    n_nodes = td["action_mask"].shape[-1]
    if env_name == "pdp":
        return (n_nodes - 1) // 2
    return n_nodes - 1 if env_name in _DEPOT_ENVS else n_nodes


def select_start_nodes(td, env, num_starts: int) -> Tensor:
    """Forced first action of rollout ``r = s*B + b``: node ``s mod num_loc``, shifted past the depot where there is
    one (utils/ops.py:139-163; the op / mtsp resampling branches are out of scope).  This is the formula the rollout
    kernel applies on the device (csrc/am_rollout*.cu)."""
    num_loc = getattr(env.generator, "num_loc", 0xFFFFFFFF)
    batch = td.shape[0]
    start_of_row = torch.arange(num_starts * batch, device=td.device) // batch
    first = start_of_row % num_loc
    return first if env.name in _NO_DEPOT_START_ENVS else first + 1
