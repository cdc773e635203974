"""Evaluation drivers on device-resident datasets, the reference's names.

Follows rl4co/tasks/eval.py:18-305: ``EvalBase.__call__`` (batches -> ``env.reset`` -> ``_inner`` -> cat,
zero-pad actions to the longest batch, report ``actions / rewards / inference_time / avg_reward``) and the
``_inner`` of ``GreedyEval``, ``AugmentationEval``, ``SamplingEval``, ``GreedyMultiStartEval`` and
``GreedyMultiStartAugmentEval``.  Differences, all on the host side of the kernels:

* the "dataloader" is a device TensorDict (or ``data.TensorDictDataset``) sliced in place, or any iterable
  of batches; nothing is collated on the host;
* rewards come from the rollout kernel (same value as ``env.get_reward(td_init, actions)`` of the reference:
  augmentation moves ``locs`` only for the policy's encoder, so the augmentation drivers re-score the tours
  on the un-augmented instance with the fused replay kernel, exactly like eval.py:137 / :293);
* the best-of-(starts x augmentations) reduction and the best-tour gather run on the device;
* ``SamplingEval``: the reference's ``multisample`` decoding fixes the first action of each of the S samples
  to a random feasible node (``sample_n_random_actions``) - here each of the S samples is a plain
  single-start sampled rollout (first action drawn from the policy); ``top_p`` / ``top_k`` are not
  implemented by the fused kernel and raise.
"""
from __future__ import annotations

import time
from typing import Iterable

import torch

from .data import TensorDictDataset
from .ops import batchify, gather_by_index, unbatchify
from .tensordict_compat import is_tensordict
from .transforms import StateAugmentation


def check_unused_kwargs(class_, kwargs):
    if len(kwargs) > 0 and not (len(kwargs) == 1 and "progress" in kwargs):
        print(f"Warning: {class_.__class__.__name__} does not use kwargs {kwargs}")


def _batches(dataloader, batch_size):
    if isinstance(dataloader, TensorDictDataset):
        return dataloader.batches(batch_size)
    if is_tensordict(dataloader):
        return TensorDictDataset(dataloader).batches(batch_size)
    return dataloader       # any iterable of TensorDict batches (e.g. a torch DataLoader)


class EvalBase:
    """eval.py:18-86."""

    name = "base"

    def __init__(self, env, progress=True, batch_size: int = 65536, **kwargs):
        check_unused_kwargs(self, kwargs)
        self.env = env
        self.progress = progress
        self.batch_size = batch_size

    def __call__(self, policy, dataloader, **kwargs):
        start = time.time()
        device = next(policy.parameters()).device
        with torch.inference_mode():
            rewards_list, actions_list = [], []
            for batch in _batches(dataloader, self.batch_size):
                td = batch.to(device)
                td = self.env.reset(td)
                actions, rewards = self._inner(policy, td, **kwargs)
                rewards_list.append(rewards)
                actions_list.append(actions)
            rewards = torch.cat(rewards_list)
            max_length = max(a.size(-1) for a in actions_list)      # pad actions with zeros (eval.py:55-63)
            actions = torch.cat([torch.nn.functional.pad(a, (0, max_length - a.size(-1))) for a in actions_list], 0)
        if device.type == "cuda":
            torch.cuda.synchronize(device)
        inference_time = time.time() - start
        if self.progress:
            print(f"Mean reward for {self.name}: {rewards.mean():.4f}")
            print(f"Time: {inference_time:.4f}s")
        return {"actions": actions.cpu(), "rewards": rewards.cpu(), "inference_time": inference_time,
                "avg_reward": rewards.cpu().mean()}

    def _inner(self, policy, td):
        raise NotImplementedError("Implement in subclass")


class GreedyEval(EvalBase):
    """eval.py:88-105: greedy decoding, single trajectory."""

    name = "greedy"

    def __init__(self, env, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)

    def _inner(self, policy, td):
        out = policy(td, self.env, decode_type="greedy", num_starts=0, return_actions=True)
        return out["actions"], out["reward"]       # = env.get_reward(td, actions), computed by the rollout kernel


class AugmentationEval(EvalBase):
    """eval.py:108-148: greedy on ``num_augment`` augmented copies, best copy per instance."""

    name = "augmentation"

    def __init__(self, env, num_augment=8, force_dihedral_8=False, feats=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        self.augmentation = StateAugmentation(num_augment=num_augment,
                                              augment_fn="dihedral8" if force_dihedral_8 else "symmetric", feats=feats)

    def _inner(self, policy, td, num_augment=None):
        if num_augment is None:
            num_augment = self.augmentation.num_augment
        td_init = td
        td_aug = self.augmentation(td)          # reset state replicated, ``locs`` (depot included) transformed
        out = policy(td_aug, self.env, decode_type="greedy", num_starts=0, return_actions=True)
        rewards = self.env.get_reward(batchify(td_init, num_augment), out["actions"])    # on the original coordinates
        rewards = unbatchify(rewards, num_augment)
        actions = unbatchify(out["actions"], num_augment)
        rewards, max_idxs = rewards.max(dim=1)
        return gather_by_index(actions, max_idxs, dim=1), rewards

    @property
    def num_augment(self):
        return self.augmentation.num_augment


class SamplingEval(EvalBase):
    """eval.py:151-201: ``samples`` sampled rollouts per instance, best one kept if ``select_best``."""

    name = "sampling"

    def __init__(self, env, samples, softmax_temp=None, select_best=True, temperature=1.0, top_p=0.0, top_k=0,
                 **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        if top_p or top_k:
            raise NotImplementedError("top_p / top_k are not implemented by the fused rollout kernel")
        self.samples = samples
        self.softmax_temp = softmax_temp
        self.temperature = temperature
        self.select_best = select_best

    def _inner(self, policy, td):
        S = self.samples
        temperature = self.softmax_temp if self.softmax_temp is not None else self.temperature
        out = policy(batchify(td, S), self.env, decode_type="sampling", temperature=temperature, return_actions=True)
        rewards, actions = out["reward"], out["actions"]
        if self.select_best:                                    # decoding.py:399-407
            rewards, max_idxs = unbatchify(rewards, S).max(dim=-1)
            actions = gather_by_index(unbatchify(actions, S), max_idxs, dim=1)
        return actions, rewards


class GreedyMultiStartEval(EvalBase):
    """eval.py:204-238: ``num_starts`` greedy multistart rollouts, best start per instance."""

    name = "multistart_greedy"

    def __init__(self, env, num_starts=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        assert num_starts is not None, "Must specify num_starts"
        self.num_starts = num_starts

    def _inner(self, policy, td):
        out = policy(td, self.env, decode_type="multistart_greedy", num_starts=self.num_starts, return_actions=True)
        rewards = unbatchify(out["reward"], self.num_starts)
        actions = unbatchify(out["actions"], self.num_starts)
        rewards, max_idxs = rewards.max(dim=1)
        return gather_by_index(actions, max_idxs, dim=1), rewards


class GreedyMultiStartAugmentEval(EvalBase):
    """eval.py:241-305: multistart greedy on augmented copies, best of starts x augmentations (POMO test)."""

    name = "multistart_greedy_augment"

    def __init__(self, env, num_starts=None, num_augment=8, force_dihedral_8=False, feats=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        assert num_starts is not None, "Must specify num_starts"
        self.num_starts = num_starts
        assert not (num_augment != 8 and force_dihedral_8), "Cannot force dihedral 8 when num_augment != 8"
        self.augmentation = StateAugmentation(num_augment=num_augment,
                                              augment_fn="dihedral8" if force_dihedral_8 else "symmetric", feats=feats)

    def _inner(self, policy, td, num_augment=None):
        if num_augment is None:
            num_augment = self.augmentation.num_augment
        td_init = td
        td_aug = self.augmentation(td)
        out = policy(td_aug, self.env, decode_type="multistart_greedy", num_starts=self.num_starts,
                     return_actions=True)
        td_rep = batchify(td_init, (num_augment, self.num_starts))
        rewards = self.env.get_reward(td_rep, out["actions"])
        rewards = unbatchify(rewards, self.num_starts * num_augment)
        actions = unbatchify(out["actions"], self.num_starts * num_augment)
        rewards, max_idxs = rewards.max(dim=1)
        return gather_by_index(actions, max_idxs, dim=1), rewards

    @property
    def num_augment(self):
        return self.augmentation.num_augment


def evaluate_policy(env, policy, dataset, method: str = "greedy", batch_size: int = 65536, **kwargs):
    """eval.py:339-420 without the automatic batch-size search and the file output."""
    methods = {
        "greedy": (GreedyEval, {}),
        "sampling": (SamplingEval, {"samples": kwargs.pop("samples", 100), "softmax_temp": kwargs.pop("softmax_temp", 1.0)}),
        "multistart_greedy": (GreedyMultiStartEval, {"num_starts": kwargs.pop("num_starts", None)}),
        "augment_dihedral_8": (AugmentationEval, {"num_augment": 8, "force_dihedral_8": True}),
        "augment": (AugmentationEval, {"num_augment": kwargs.pop("num_augment", 8)}),
        "multistart_greedy_augment_dihedral_8": (GreedyMultiStartAugmentEval,
                                                 {"num_augment": 8, "force_dihedral_8": True,
                                                  "num_starts": kwargs.get("num_starts_aug", None)}),
        "multistart_greedy_augment": (GreedyMultiStartAugmentEval, {"num_augment": kwargs.pop("num_augment_ms", 8),
                                                                    "num_starts": kwargs.get("num_starts_aug", None)}),
    }
    assert method in methods, "Method {} not found".format(method)
    cls, kw = methods[method]
    if "num_starts" in kw and kw["num_starts"] is None:
        kw["num_starts"] = env.get_num_starts(env.reset(dataset[:1] if is_tensordict(dataset) else dataset[0:1]))
    kwargs.pop("num_starts_aug", None)
    eval_fn = cls(env, batch_size=batch_size, **kw, **kwargs)
    return eval_fn(policy, dataset)
