"""Minimal TensorDict-compatible container.

The reference passes every state through ``tensordict.TensorDict`` (third party, not
installed on the GPU box).  The rollout path only needs a dict of tensors with a leading
batch shape plus the handful of methods listed in SURVEY.md Appendix B, so this module
provides exactly that.  If the real ``tensordict`` package is importable it is used
instead, so a td produced by either side is consumable by the other.

Reference call sites that define the needed surface:
  rl4co/utils/ops.py:11-62 (expand/contiguous/view/permute on batch dims),
  rl4co/envs/common/base.py:135-143 (reset merges into the input td),
  rl4co/envs/routing/tsp/env.py:83-93 (td.update in place).
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, Optional

import torch

try:  # pragma: no cover - real package absent in this image
    from tensordict import TensorDict as _RealTensorDict  # type: ignore

    HAVE_REAL_TENSORDICT = True
except Exception:  # noqa: BLE001
    _RealTensorDict = None
    HAVE_REAL_TENSORDICT = False


def _as_size(batch_size) -> torch.Size:
    if batch_size is None:
        return torch.Size([])
    if isinstance(batch_size, int):
        return torch.Size([batch_size])
    return torch.Size(list(batch_size))


class CompatTensorDict:
    """Dict of tensors sharing leading ``batch_size`` dims."""

    def __init__(self, source: Optional[Dict[str, Any]] = None, batch_size=None,
                 device=None, _run_checks: bool = False, **_):
        self._d: Dict[str, Any] = {}
        self._batch_size = _as_size(batch_size)
        self._device = torch.device(device) if device is not None else None
        if source is not None:
            items = source.items() if hasattr(source, "items") else source
            for k, v in items:
                self._d[k] = v
        if batch_size is None and self._d:
            # infer: common leading dims is ambiguous; the reference always passes batch_size
            first = next(iter(self._d.values()))
            if isinstance(first, torch.Tensor) and first.dim() > 0:
                self._batch_size = torch.Size([first.shape[0]])

    # ---- basic container protocol -------------------------------------------------
    @property
    def batch_size(self) -> torch.Size:
        return self._batch_size

    @batch_size.setter
    def batch_size(self, value):
        self._batch_size = _as_size(value)

    @property
    def shape(self) -> torch.Size:
        return self._batch_size

    @property
    def device(self):
        if self._device is not None:
            return self._device
        for v in self._d.values():
            if isinstance(v, torch.Tensor):
                return v.device
        return None

    def dim(self) -> int:
        return len(self._batch_size)

    ndim = property(dim)

    def size(self, i: Optional[int] = None):
        return self._batch_size if i is None else self._batch_size[i]

    def numel(self) -> int:
        n = 1
        for s in self._batch_size:
            n *= s
        return n

    def keys(self, *a, **k):
        return self._d.keys()

    def values(self, *a, **k):
        return self._d.values()

    def items(self, *a, **k):
        return self._d.items()

    def __contains__(self, key) -> bool:
        return key in self._d

    def __iter__(self):
        return iter(self._d)

    def __len__(self) -> int:
        return self._batch_size[0] if len(self._batch_size) else 0

    def is_empty(self) -> bool:
        return len(self._d) == 0

    def _map(self, fn, batch_size) -> "CompatTensorDict":
        out = CompatTensorDict({}, batch_size=batch_size, device=self._device)
        for k, v in self._d.items():
            out._d[k] = fn(v) if isinstance(v, (torch.Tensor, CompatTensorDict)) else v
        return out

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return self._d[idx]
        probe = torch.empty(self._batch_size, device="meta")[idx]
        return self._map(lambda v: v[idx], probe.shape)

    def __setitem__(self, key, value):
        if isinstance(key, str):
            self._d[key] = value
        else:
            for k, v in self._d.items():
                v[key] = value[k] if isinstance(value, (CompatTensorDict, dict)) else value

    def get(self, key, default=None):
        return self._d.get(key, default)

    def set(self, key, value, inplace: bool = False):
        self._d[key] = value
        return self

    def pop(self, key, *default):
        return self._d.pop(key, *default)

    def update(self, other, **_):
        items = other.items() if hasattr(other, "items") else other
        for k, v in items:
            self._d[k] = v
        return self

    def select(self, *keys, strict: bool = True):
        out = CompatTensorDict({}, batch_size=self._batch_size, device=self._device)
        for k in keys:
            if k in self._d:
                out._d[k] = self._d[k]
            elif strict:
                raise KeyError(k)
        return out

    def exclude(self, *keys):
        out = CompatTensorDict({}, batch_size=self._batch_size, device=self._device)
        for k, v in self._d.items():
            if k not in keys:
                out._d[k] = v
        return out

    def clone(self, recurse: bool = True):
        return self._map(lambda v: v.clone() if recurse else v, self._batch_size)

    def copy(self):
        return self.clone(recurse=False)

    def to(self, device, **kw):
        if device is None:
            return self
        out = self._map(lambda v: v.to(device, **kw), self._batch_size)
        out._device = torch.device(device) if not isinstance(device, torch.dtype) else self._device
        return out

    def cpu(self):
        return self.to("cpu")

    def cuda(self, index=None):
        return self.to("cuda" if index is None else f"cuda:{index}")

    def detach(self):
        return self._map(lambda v: v.detach(), self._batch_size)

    def to_dict(self):
        return dict(self._d)

    # ---- batch-dim reshapes used by batchify / unbatchify ------------------------
    def _nb(self) -> int:
        return len(self._batch_size)

    def expand(self, *shape):
        shape = _as_size(shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape)
        nb = self._nb()
        return self._map(lambda v: v.expand(*shape, *v.shape[nb:]), shape)

    def contiguous(self):
        return self._map(lambda v: v.contiguous(), self._batch_size)

    def view(self, *shape):
        shape = _as_size(shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape)
        nb = self._nb()
        return self._map(lambda v: v.view(*shape, *v.shape[nb:]), shape)

    def reshape(self, *shape):
        shape = _as_size(shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape)
        nb = self._nb()
        return self._map(lambda v: v.reshape(*shape, *v.shape[nb:]), shape)

    def permute(self, *dims):
        dims = tuple(dims[0]) if len(dims) == 1 and not isinstance(dims[0], int) else tuple(dims)
        nb = self._nb()
        bdims = dims[:nb]
        new_bs = torch.Size([self._batch_size[d] for d in bdims])
        return self._map(
            lambda v: v.permute(*bdims, *range(nb, v.dim())), new_bs
        )

    def gather(self, dim, index):
        nb = self._nb()

        def g(v):
            idx = index.view(index.shape + (1,) * (v.dim() - index.dim()))
            idx = idx.expand(*index.shape[: dim + 1], *v.shape[dim + 1:]) if v.dim() > index.dim() else idx
            return v.gather(dim, idx)

        return self._map(g, index.shape[:nb])

    def squeeze(self, dim):
        bs = list(self._batch_size)
        if bs[dim] == 1:
            bs.pop(dim)
            return self._map(lambda v: v.squeeze(dim), bs)
        return self

    def __repr__(self) -> str:
        body = ", ".join(
            f"{k}: {tuple(v.shape) if isinstance(v, torch.Tensor) else type(v).__name__}"
            for k, v in self._d.items()
        )
        return f"TensorDict({{{body}}}, batch_size={list(self._batch_size)})"


TensorDict = _RealTensorDict if HAVE_REAL_TENSORDICT else CompatTensorDict


def is_tensordict(x) -> bool:
    return isinstance(x, CompatTensorDict) or (
        HAVE_REAL_TENSORDICT and isinstance(x, _RealTensorDict)
    )
