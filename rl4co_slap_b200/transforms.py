"""State augmentation for POMO-style evaluation, the reference's names.

Follows rl4co/data/transforms.py:15-150: ``dihedral_8_augmentation`` (the eight symmetries of the
unit square, augmentation-major ``[8*B, N, 2]``), ``symmetric_augmentation`` (random rotation +
reflection about (0.5, 0.5), first 1/num_augment of the batch kept as is) and ``StateAugmentation``
(``batchify`` the whole td, then transform the listed features).  Pure elementwise work on the
device the data lives on: nothing is copied to the host.
"""
from __future__ import annotations

import math
from typing import Callable, Union

import torch
from torch import Tensor

from .ops import batchify


def dihedral_8_augmentation(xy: Tensor) -> Tensor:
    """transforms.py:15-37: [B, N, 2] -> [8*B, N, 2], blocks (x,y) (1-x,y) (x,1-y) (1-x,1-y) (y,x) (1-y,x)
    (y,1-x) (1-y,1-x)."""
    x, y = xy[..., 0:1], xy[..., 1:2]
    zs = ((x, y), (1 - x, y), (x, 1 - y), (1 - x, 1 - y), (y, x), (1 - y, x), (y, 1 - x), (1 - y, 1 - x))
    return torch.cat([torch.cat(z, dim=2) for z in zs], dim=0)


def dihedral_8_augmentation_wrapper(xy: Tensor, reduce: bool = True, *args, **kw) -> Tensor:
    """transforms.py:40-47: the input is already batchified x8; transform its first eighth."""
    xy = xy[: xy.shape[0] // 8, ...] if reduce else xy
    return dihedral_8_augmentation(xy)


def symmetric_transform(x: Tensor, y: Tensor, phi: Tensor, offset: float = 0.5) -> Tensor:
    """transforms.py:50-71: rotate by phi about (offset, offset); phi > 2*pi additionally swaps the axes."""
    x, y = x - offset, y - offset
    x_prime = torch.cos(phi) * x - torch.sin(phi) * y
    y_prime = torch.sin(phi) * x + torch.cos(phi) * y
    mask = phi > 2 * math.pi
    xy = torch.cat((x_prime, y_prime), dim=-1)
    xy = torch.where(mask, xy.flip(-1), xy)
    return xy + offset


def symmetric_augmentation(xy: Tensor, num_augment: int = 8, first_augment: bool = False) -> Tensor:
    """transforms.py:74-89: phi ~ U(0, 4*pi) per (augmented) instance, 0 for the first copy."""
    phi = torch.rand(xy.shape[0], device=xy.device) * 4 * math.pi
    if not first_augment:
        phi[: xy.shape[0] // num_augment] = 0.0
    x, y = xy[..., [0]], xy[..., [1]]
    return symmetric_transform(x, y, phi[:, None, None])


def min_max_normalize(x: Tensor) -> Tensor:
    return (x - x.min()) / (x.max() - x.min())


def get_augment_function(augment_fn: Union[str, Callable]):
    if callable(augment_fn):
        return augment_fn
    if augment_fn == "dihedral8":
        return dihedral_8_augmentation_wrapper
    if augment_fn == "symmetric":
        return symmetric_augmentation
    raise ValueError(f"Unknown augment_fn: {augment_fn}. Available options: 'symmetric', 'dihedral8' or a custom callable")


class StateAugmentation:
    """transforms.py:104-150.  ``td_aug = batchify(td, num_augment)`` then every feature in ``feats``
    (default ``["locs"]``) is replaced by its augmentation; copy a of instance b sits at a*B + b."""

    def __init__(self, num_augment: int = 8, augment_fn: Union[str, Callable] = "symmetric",
                 first_aug_identity: bool = True, normalize: bool = False, feats: list = None):
        self.augmentation = get_augment_function(augment_fn)
        assert not (self.augmentation == dihedral_8_augmentation_wrapper and num_augment != 8), \
            "When using the `dihedral8` augmentation function, then num_augment must be 8"
        self.feats = ["locs"] if feats is None else feats
        self.num_augment = num_augment
        self.normalize = normalize
        self.first_aug_identity = first_aug_identity

    def __call__(self, td):
        td_aug = batchify(td, self.num_augment)
        idx = list(td.batch_size)           # the reference indexes with list(td.size()) (transforms.py:141,146)
        for feat in self.feats:
            init = None
            if not self.first_aug_identity:
                init = td_aug[feat][idx, 0].clone()
            aug_feat = self.augmentation(td_aug[feat], self.num_augment)
            if self.normalize:
                aug_feat = min_max_normalize(aug_feat)
            if init is not None:
                aug_feat[idx, 0] = init
            td_aug[feat] = aug_feat
        return td_aug
