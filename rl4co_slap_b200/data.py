"""Datasets for the rollout path: Kool-format ``.npz`` IO and on-device generation.

Follows rl4co/data/utils.py:11-30 (``load_npz_to_tensordict`` / ``save_tensordict_to_npz``: a flat dict
of arrays, batch size = leading dim of the first key), rl4co/envs/routing/cvrp/env.py:192-199
(``CVRPEnv.load_data``: demand / capacity), rl4co/data/generate_data.py:40-83,324-335 (keys ``locs``,
``depot``, ``demand``, ``capacity``; val seed 4321, test seed 1234) and rl4co/data/dataset.py
(``TensorDictDataset.add_key`` used by ``RolloutBaseline.wrap_dataset``).

A dataset here is one device-resident TensorDict: the eval drivers and the rollout baseline slice it
in place instead of collating host batches (1 Mi TSP100 instances are 839 MB: one H2D copy, not one per
epoch).  ``generate_dataset`` samples with the env's own (device) generator, so no host copy exists at all.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

from .tensordict_compat import TensorDict


def check_extension(filename: str, extension: str = ".npz") -> str:
    """data/utils.py:33-37."""
    if os.path.splitext(filename)[1] != extension:
        return filename + extension
    return filename


def load_npz_to_tensordict(filename: str, device=None) -> TensorDict:
    """data/utils.py:11-19; ``device``: optional target (one pinned, non-blocking copy per key)."""
    x = np.load(filename)
    x_dict = {k: torch.from_numpy(np.ascontiguousarray(x[k])) for k in x.files}
    batch_size = next(iter(x_dict.values())).shape[0]
    td = TensorDict(x_dict, batch_size=[batch_size])
    return td if device is None else to_device(td, device)


def save_tensordict_to_npz(tensordict, filename: str, compress: bool = False) -> None:
    """data/utils.py:22-30 (tensors are brought to the host first)."""
    x_dict = {k: v.detach().cpu().numpy() for k, v in tensordict.items()}
    (np.savez_compressed if compress else np.savez)(filename, **x_dict)


def to_device(td, device):
    """Host TensorDict -> device through pinned staging buffers (asynchronous copies on the current stream)."""
    device = torch.device(device)
    if device.type != "cuda":
        return td.to(device)
    out = {}
    for k, v in td.items():
        if v.is_cuda:
            out[k] = v.to(device)
            continue
        try:
            v = v.pin_memory()
        except RuntimeError:            # no CUDA context for pinning: plain copy
            pass
        out[k] = v.to(device, non_blocking=True)
    return TensorDict(out, batch_size=td.batch_size)


def load_data(env_name: str, fpath: str, device=None) -> TensorDict:
    """``Env.load_data`` of the three environments: TSP / SLAP load the file as is
    (envs/common/base.py:270-273), CVRP normalises demand by capacity (cvrp/env.py:192-199)."""
    td = load_npz_to_tensordict(fpath)
    if env_name == "cvrp":
        td.set("demand", td["demand"] / td["capacity"][:, None])
    return td if device is None else to_device(td, device)


def _retarget(gen, dev):
    """Point a generator and its device-side samplers at ``dev``; returns the undo list."""
    undo = []
    for obj in [gen] + [v for v in vars(gen).values() if not isinstance(v, torch.Tensor)]:
        if hasattr(obj, "device") and not callable(getattr(obj, "device")):
            try:
                undo.append((obj, obj.device))
                obj.device = dev
            except AttributeError:
                undo.pop()
    return undo


def generate_dataset(env, dataset_size: int, seed: Optional[int] = None, device=None) -> TensorDict:
    """``env.dataset(batch_size=[dataset_size])`` of the reference with no file (envs/common/base.py:236-243):
    instances from the env's generator, sampled directly on ``device`` (default: the env's device)."""
    gen = env.generator
    dev = torch.device(device) if device is not None else getattr(env, "device", None)
    if seed is not None:
        torch.manual_seed(seed)
    undo = _retarget(gen, dev) if dev is not None else []
    try:
        td = gen(batch_size=[dataset_size])
    finally:
        for obj, old in undo:
            obj.device = old
    return td if dev is None or td.device == dev else td.to(dev)


def generate_default_datasets_spec():
    """generate_data.py:324-335: (name, seed) of the default validation / test sets."""
    return (("val", 4321), ("test", 1234))


def generate_npz(problem: str, dataset_size: int, graph_size: int, seed: int) -> dict:
    """generate_data.py:40-83: numpy RNG, the reference's draw order, for files that must match the
    reference's published val / test sets byte for byte."""
    from .generators import CAPACITIES
    np.random.seed(seed)
    if problem == "tsp":
        return {"locs": np.random.uniform(size=(dataset_size, graph_size, 2)).astype(np.float32)}
    if problem in ("cvrp", "vrp"):
        return {
            "depot": np.random.uniform(size=(dataset_size, 2)).astype(np.float32),
            "locs": np.random.uniform(size=(dataset_size, graph_size, 2)).astype(np.float32),
            "demand": np.random.randint(1, 10, size=(dataset_size, graph_size)).astype(np.float32),
            "capacity": np.full(dataset_size, CAPACITIES[graph_size]).astype(np.float32),
        }
    raise ValueError(f"no npz generator for problem {problem!r} on this path")


class TensorDictDataset:
    """data/dataset.py ``TensorDictDataset``: a thin, sliceable view of one TensorDict; ``add_key`` is what
    ``RolloutBaseline.wrap_dataset`` uses to attach the baseline rewards as ``"extra"``."""

    def __init__(self, td):
        self.data = td
        self.data_len = td.batch_size[0]

    def __len__(self):
        return self.data_len

    def __getitem__(self, idx):
        return self.data[idx]

    def add_key(self, key, value):
        self.data[key] = value
        return self

    def batches(self, batch_size: int):
        """Contiguous device slices (views, no collate)."""
        for lo in range(0, self.data_len, batch_size):
            yield self.data[lo:min(self.data_len, lo + batch_size)]

    @staticmethod
    def collate_fn(batch):
        return batch
