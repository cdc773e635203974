"""Evaluation drivers on device-resident datasets.

Every driver of rl4co/tasks/eval.py:88-305 is the same pipeline with different settings - how many augmented copies,
how many starts or samples, greedy or sampled - so here ONE routine (:func:`run_plan`) executes a declarative
:class:`Plan` and the reference's driver names are thin declarations of their plan:

    instances -> [augment, csrc/pomo_epilogue.cu] -> policy (K3 + K2) -> [re-score on the un-augmented instance,
    fused replay kernel] -> best of copies x starts + winning tours (rl4co_pomo_best, one launch)

Kept from the reference: class names, constructor keywords, ``driver(policy, dataloader)`` returning
``actions / rewards / inference_time / avg_reward`` (eval.py:30-83), zero padding of tours to the longest batch, and
that augmented tours are scored on the ORIGINAL coordinates (eval.py:137, :293).  Different: the "dataloader" is a
device TensorDict sliced in place (or any iterable of batches), nothing is collated on the host, and ``top_p`` /
``top_k`` raise because the fused kernel does not implement them.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Optional

import torch

from .data import TensorDictDataset
from .ops import batchify
from .pomo import best_of_rollouts
from .tensordict_compat import is_tensordict
from .transforms import StateAugmentation


@dataclass
class Plan:
    """What one evaluation method does per batch."""
    decode_type: str = "greedy"
    num_starts: int = 0                 # POMO starts (first action forced) per instance or augmented copy
    num_samples: int = 0                # independent sampled rollouts per instance (multisample decoding)
    augmentation: Optional[StateAugmentation] = None
    temperature: Optional[float] = None
    select_best: bool = True


def run_plan(policy, env, td, plan: Plan):
    """-> (actions, rewards): ``[B, T]`` / ``[B]`` when the best rollout per instance is selected, else every rollout."""
    instances = int(td.batch_size[0])
    n_aug = plan.augmentation.num_augment if plan.augmentation is not None else 0
    repeats = plan.num_starts or plan.num_samples
    td_in = plan.augmentation(td) if n_aug > 1 else td
    kwargs = {"decode_type": plan.decode_type, "return_actions": True, "num_starts": repeats}
    if plan.num_samples:
        kwargs["multisample"] = True
    if plan.temperature is not None:
        kwargs["temperature"] = plan.temperature
    out = policy(td_in, env, **kwargs)
    actions, rewards = out["actions"], out["reward"]
    if n_aug > 1:   # the policy saw transformed coordinates; score its tours on the instance itself
        rewards = env.get_reward(batchify(td, (n_aug, repeats)), actions)
    if not plan.select_best or (n_aug <= 1 and repeats <= 1):
        return actions, rewards
    best = best_of_rollouts(rewards, actions, instances, n_aug, repeats)
    return best["best_aug_actions"], best["max_aug_reward"]


def _batches(dataloader, batch_size):
    if isinstance(dataloader, TensorDictDataset):
        return dataloader.batches(batch_size)
    if is_tensordict(dataloader):
        return TensorDictDataset(dataloader).batches(batch_size)
    return dataloader       # any iterable of TensorDict batches


def check_unused_kwargs(class_, kwargs):
    if len(kwargs) > 0 and not (len(kwargs) == 1 and "progress" in kwargs):
        print(f"Warning: {class_.__class__.__name__} does not use kwargs {kwargs}")


class EvalBase:
    """eval.py:18-86: run a method over a dataset, collect per-instance tours and rewards."""

    name = "base"

    def __init__(self, env, progress=True, batch_size: int = 65536, **kwargs):
        check_unused_kwargs(self, kwargs)
        self.env, self.progress, self.batch_size = env, progress, batch_size

    def plan(self) -> Plan:
        raise NotImplementedError("Implement in subclass")

    def _inner(self, policy, td, **kwargs):
        return run_plan(policy, self.env, td, self.plan())

    def __call__(self, policy, dataloader, **kwargs):
        tic = time.time()
        device = next(policy.parameters()).device
        tours, scores = [], []
        with torch.inference_mode():
            for batch in _batches(dataloader, self.batch_size):
                actions, rewards = self._inner(policy, self.env.reset(batch.to(device)), **kwargs)
                tours.append(actions)
                scores.append(rewards)
            width = max(a.shape[-1] for a in tours)             # routes differ in length: zero pad (eval.py:55-63)
            actions = torch.cat([torch.nn.functional.pad(a, (0, width - a.shape[-1])) for a in tours])
            rewards = torch.cat(scores)
        if device.type == "cuda":
            torch.cuda.synchronize(device)
        elapsed = time.time() - tic
        if self.progress:
            print(f"Mean reward for {self.name}: {rewards.mean():.4f}")
            print(f"Time: {elapsed:.4f}s")
        rewards = rewards.cpu()
        return {"actions": actions.cpu(), "rewards": rewards, "inference_time": elapsed, "avg_reward": rewards.mean()}


def _augmenter(num_augment, force_dihedral_8, feats):
    assert not (num_augment != 8 and force_dihedral_8), "Cannot force dihedral 8 when num_augment != 8"
    return StateAugmentation(num_augment=num_augment, augment_fn="dihedral8" if force_dihedral_8 else "symmetric",
                             feats=feats)


class GreedyEval(EvalBase):
    """eval.py:88-105: one greedy rollout per instance."""
    name = "greedy"

    def __init__(self, env, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)

    def plan(self):
        return Plan("greedy")


class AugmentationEval(EvalBase):
    """eval.py:108-148: greedy on ``num_augment`` augmented copies, best copy per instance."""
    name = "augmentation"

    def __init__(self, env, num_augment=8, force_dihedral_8=False, feats=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        self.augmentation = _augmenter(num_augment, force_dihedral_8, feats)

    @property
    def num_augment(self):
        return self.augmentation.num_augment

    def plan(self):
        return Plan("greedy", augmentation=self.augmentation)


class SamplingEval(EvalBase):
    """eval.py:151-201: ``samples`` sampled rollouts per instance through the policy's ``multisample`` decoding (the td
    is tiled and EVERY action, the first included, is drawn from the policy - rl4co/utils/decoding.py:306-308; the
    ``select_start_nodes_fn`` the reference passes is only consulted by multistart decoding, :286-291)."""
    name = "sampling"

    def __init__(self, env, samples, softmax_temp=None, select_best=True, temperature=1.0, top_p=0.0, top_k=0, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        if top_p or top_k:
            raise NotImplementedError("top_p / top_k are not implemented by the fused rollout kernel")
        self.samples, self.select_best = samples, select_best
        self.softmax_temp, self.temperature = softmax_temp, temperature

    def plan(self):
        temperature = self.softmax_temp if self.softmax_temp is not None else self.temperature
        return Plan("sampling", num_samples=self.samples, temperature=temperature, select_best=self.select_best)


class GreedyMultiStartEval(EvalBase):
    """eval.py:204-238: greedy from ``num_starts`` forced start nodes, best start per instance."""
    name = "multistart_greedy"

    def __init__(self, env, num_starts=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        assert num_starts is not None, "Must specify num_starts"
        self.num_starts = num_starts

    def plan(self):
        return Plan("multistart_greedy", num_starts=self.num_starts)


class GreedyMultiStartAugmentEval(EvalBase):
    """eval.py:241-305: multistart greedy on augmented copies, best of copies x starts (the POMO test protocol)."""
    name = "multistart_greedy_augment"

    def __init__(self, env, num_starts=None, num_augment=8, force_dihedral_8=False, feats=None, **kwargs):
        super().__init__(env, kwargs.pop("progress", True), **kwargs)
        assert num_starts is not None, "Must specify num_starts"
        self.num_starts = num_starts
        self.augmentation = _augmenter(num_augment, force_dihedral_8, feats)

    @property
    def num_augment(self):
        return self.augmentation.num_augment

    def plan(self):
        return Plan("multistart_greedy", num_starts=self.num_starts, augmentation=self.augmentation)


# method name -> (driver, its default keywords); eval.py:352-384.  ``num_starts: None`` = one per candidate start node.
_METHODS = {
    "greedy": (GreedyEval, {}),
    "sampling": (SamplingEval, {"samples": 1280, "softmax_temp": 1.0}),
    "multistart_greedy": (GreedyMultiStartEval, {"num_starts": None}),
    "augment_dihedral_8": (AugmentationEval, {"num_augment": 8, "force_dihedral_8": True}),
    "augment": (AugmentationEval, {"num_augment": 8}),
    "multistart_greedy_augment_dihedral_8": (GreedyMultiStartAugmentEval,
                                             {"num_augment": 8, "force_dihedral_8": True, "num_starts": None}),
    "multistart_greedy_augment": (GreedyMultiStartAugmentEval, {"num_augment": 8, "num_starts": None}),
}


def evaluate_policy(env, policy, dataset, method: str = "greedy", batch_size: int = 65536, **kwargs):
    """eval.py:339-420 without the automatic batch-size search and the file output: pick the method, start from ITS
    defaults, then let the caller's keywords override them (eval.py:385-392)."""
    assert method in _METHODS, "Method {} not found".format(method)
    driver, defaults = _METHODS[method]
    settings = {**defaults, **kwargs}
    if "num_starts" in defaults and settings.get("num_starts") is None:
        first = dataset[:1] if is_tensordict(dataset) else dataset[0:1]
        settings["num_starts"] = env.get_num_starts(env.reset(first))
    return driver(env, batch_size=batch_size, **settings)(policy, dataset)
