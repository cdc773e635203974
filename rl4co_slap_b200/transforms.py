"""Instance augmentation on the way INTO the encoder batch - kernel backed.

What the reference does with torch ops on an already 8-fold tiled tensor (rl4co/data/transforms.py:15-150), this
module does with one launch that reads each coordinate once and writes its augmented copies
(csrc/pomo_epilogue.cu: ``rl4co_dihedral8_augment``, ``rl4co_symmetric_augment``).  The names callers use are the
reference's - ``StateAugmentation``, ``dihedral_8_augmentation``, ``symmetric_augmentation``,
``get_augment_function`` - and so is the layout: copy ``a`` of instance ``b`` is row ``a*B + b``.

CUDA tensors only: there is no host arithmetic path (the library raises on CPU tensors).
"""
from __future__ import annotations

import math
from typing import Callable, Optional, Union

import torch
from torch import Tensor

from . import _lib
from .ops import batchify

_TWO_PI = 2.0 * math.pi


def _xy_rows(xy: Tensor, who: str) -> Tensor:
    if xy.dim() != 3 or xy.shape[-1] != 2:
        raise ValueError(f"{who}: expected coordinates [rows, nodes, 2], got {tuple(xy.shape)}")
    return xy.to(torch.float32).contiguous()


def dihedral_8_augmentation(xy: Tensor) -> Tensor:
    """[B, N, 2] -> [8*B, N, 2]: the eight symmetries of the unit square in the reference's copy order
    (transforms.py:15-37), written straight from the un-tiled input."""
    xy = _xy_rows(xy, "dihedral_8_augmentation")
    rows, nodes, _ = xy.shape
    out = torch.empty(8 * rows, nodes, 2, dtype=torch.float32, device=xy.device)
    _lib.call("rl4co_dihedral8_augment", rows, nodes, _lib.ptr(xy), _lib.ptr(out), _lib.stream_ptr(xy.device))
    return out


def symmetric_transform(xy: Tensor, phi: Tensor) -> Tensor:
    """Rotate every row's points by that row's ``phi`` about the centre of the unit square; ``phi > 2*pi`` also
    mirrors (swaps the coordinates) - transforms.py:50-71.  ``xy [rows, N, 2]``, ``phi [rows]``."""
    xy = _xy_rows(xy, "symmetric_transform")
    rows, nodes, _ = xy.shape
    phi = phi.reshape(-1).to(device=xy.device, dtype=torch.float32).contiguous()
    if phi.numel() != rows:
        raise ValueError(f"symmetric_transform: {rows} rows but {phi.numel()} angles")
    out = torch.empty_like(xy)
    _lib.call("rl4co_symmetric_augment", rows, nodes, _lib.ptr(xy), _lib.ptr(phi), _lib.ptr(out),
              _lib.stream_ptr(xy.device))
    return out


def symmetric_augmentation(xy: Tensor, num_augment: int = 8, first_augment: bool = False) -> Tensor:
    """Random symmetric augmentation of an already tiled batch ``[A*B, N, 2]`` (transforms.py:74-89): one angle
    ``~ U(0, 4*pi)`` per row - the upper half of that range mirrors - and angle 0 (identity) for the first copy of
    every instance unless ``first_augment``."""
    rows = xy.shape[0]
    phi = torch.rand(rows, device=xy.device) * (2.0 * _TWO_PI)
    if not first_augment:
        phi[: rows // num_augment] = 0.0
    return symmetric_transform(xy, phi)


def _dihedral8_of_tiled(tiled_xy: Tensor, num_augment: int = 8, *unused, **unused_kw) -> Tensor:
    """The reference's wrapper contract (transforms.py:40-47): handed the 8-fold tiled feature, use its first eighth."""
    return dihedral_8_augmentation(tiled_xy[: tiled_xy.shape[0] // 8])


# the reference's name for the same callable (get_augment_function returns it; StateAugmentation compares against it)
dihedral_8_augmentation_wrapper = _dihedral8_of_tiled

_AUGMENTERS = {"dihedral8": _dihedral8_of_tiled, "symmetric": symmetric_augmentation}


def get_augment_function(augment_fn: Union[str, Callable]) -> Callable:
    """transforms.py:92-101: a name from the table above, or any callable ``(tiled_feature, num_augment) -> feature``."""
    if callable(augment_fn):
        return augment_fn
    try:
        return _AUGMENTERS[augment_fn]
    except KeyError:
        raise ValueError(f"Unknown augment_fn: {augment_fn}. Available options: 'symmetric', 'dihedral8' or a custom "
                         "callable") from None


def min_max_normalize(x: Tensor) -> Tensor:
    lo, hi = torch.aminmax(x)
    return (x - lo) / (hi - lo)


class StateAugmentation:
    """``td -> td`` with the batch tiled ``num_augment`` times and the listed features (default ``locs``) replaced by
    their augmented copies (transforms.py:104-150).  Everything else in the td - demands, capacities, masks - is the
    plain tiling the rollout kernel expects for augmented copies of the same instance."""

    def __init__(self, num_augment: int = 8, augment_fn: Union[str, Callable] = "symmetric",
                 first_aug_identity: bool = True, normalize: bool = False, feats: Optional[list] = None):
        self.augmentation = get_augment_function(augment_fn)
        assert not (self.augmentation is _dihedral8_of_tiled and num_augment != 8), \
            "When using the `dihedral8` augmentation function, then num_augment must be 8"
        self.num_augment = num_augment
        self.first_aug_identity = first_aug_identity
        self.normalize = normalize
        self.feats = list(feats) if feats is not None else ["locs"]

    def __call__(self, td):
        # dihedral-8 is written straight from the un-tiled coordinates, so those are left out of the tiling
        direct = self.augmentation is _dihedral8_of_tiled
        tiled = batchify(td.exclude(*self.feats) if direct else td, self.num_augment)
        # the reference keeps (with first_aug_identity=False) node 0 of ONE row - the row whose index equals the
        # batch size (it indexes with list(td.size()), transforms.py:141-147); reproduced as is
        rows = int(td.batch_size[0])
        for name in self.feats:
            keep = None if self.first_aug_identity else td[name][0, 0].clone()    # tiled row `rows` = instance 0
            if direct:
                after = dihedral_8_augmentation(td[name])
            else:
                after = self.augmentation(tiled[name], self.num_augment)
            if self.normalize:
                after = min_max_normalize(after)
            if keep is not None:
                after[rows, 0] = keep
            tiled[name] = after
        return tiled
