"""POMO around the rollout kernels: multistart rollouts of (optionally augmented) instances and the epilogue that
picks the best start / best augmented copy - one launch of ``rl4co_pomo_best`` (csrc/pomo_epilogue.cu) instead of the
reference's unbatchify / max / gather chain.

Reference: rl4co/models/zoo/pomo/model.py:20-144 (``POMO.__init__`` arguments, ``shared_step`` output keys),
rl4co/utils/decoding.py:399-407 (``_select_best``), rl4co/models/rl/reinforce/baselines.py:57-61 (shared baseline).
The Lightning plumbing (logging, optimiser configuration, trainer hooks) is out of scope; ``shared_step`` takes the
optimiser as an argument in the train phase and returns the reference's ``out`` dictionary.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from .policy import POMOPolicy
from .transforms import StateAugmentation


def best_of_rollouts(reward: torch.Tensor, actions: Optional[torch.Tensor], batch: int, n_aug: int = 0,
                     n_start: int = 0) -> dict:
    """Segmented max over starts and over augmented copies with the winning tours.

    ``reward [R]`` / ``actions [R, T]`` hold ``R = S*A*B`` rollouts in policy output order - rollout
    ``(s*A + a)*B + b`` is start ``s`` of copy ``a`` of instance ``b`` (``A = max(n_aug, 1)``, ``S = max(n_start, 1)``).
    Returns ``max_reward [B,A]``, ``best_start [B,A]``, ``best_multistart_actions [B,A,T]``, ``max_aug_reward [B]``,
    ``best_aug [B]``, ``best_aug_actions [B,T]`` (tour outputs only when ``actions`` is given); ties go to the
    lowest index like ``torch.max``."""
    if not reward.is_cuda:
        raise RuntimeError("best_of_rollouts needs CUDA tensors (no CPU fallback)")
    A, S = max(int(n_aug), 1), max(int(n_start), 1)
    if reward.numel() != A * S * batch:
        raise ValueError(f"best_of_rollouts: {reward.numel()} rewards for {batch} instances x {A} copies x {S} starts")
    dev = reward.device
    reward = reward.reshape(-1).to(torch.float32).contiguous()
    out = {"max_reward": torch.empty(batch, A, dtype=torch.float32, device=dev),
           "best_start": torch.empty(batch, A, dtype=torch.int64, device=dev),
           "max_aug_reward": torch.empty(batch, dtype=torch.float32, device=dev),
           "best_aug": torch.empty(batch, dtype=torch.int64, device=dev)}
    T = 0
    if actions is not None:
        actions = actions.reshape(A * S * batch, -1).contiguous()
        T = actions.shape[1]
        out["best_multistart_actions"] = torch.empty(batch, A, T, dtype=torch.int64, device=dev)
        out["best_aug_actions"] = torch.empty(batch, T, dtype=torch.int64, device=dev)
    _lib.call("rl4co_pomo_best", batch, A, S, T, _lib.ptr(reward), _lib.ptr(actions, torch.int64) if actions is not None else None,
              _lib.ptr(out["max_reward"]), _lib.ptr(out["best_start"]), _lib.ptr(out.get("best_multistart_actions")),
              _lib.ptr(out["max_aug_reward"]), _lib.ptr(out["best_aug"]), _lib.ptr(out.get("best_aug_actions")),
              _lib.stream_ptr(dev))
    return out


class POMO:
    """Policy Optimization with Multiple Optima (Kwon et al.) on the fused rollout path.

    Constructor arguments follow pomo/model.py:36-56: ``num_augment`` copies per instance at validation / test time
    (``augment_fn`` "dihedral8" or "symmetric"), ``num_starts`` (default: one per candidate start node), shared
    baseline.  ``shared_step(batch, batch_idx, phase)`` mirrors pomo/model.py:87-144 and returns the dictionary the
    reference hands to its metric logger."""

    def __init__(self, env, policy=None, policy_kwargs: Optional[dict] = None, baseline: str = "shared",
                 num_augment: int = 8, augment_fn="dihedral8", first_aug_identity: bool = True, feats: Optional[list] = None,
                 num_starts: Optional[int] = None, **unused):
        assert baseline == "shared", "POMO only supports shared baseline"
        self.env = env
        self.policy = policy if policy is not None else POMOPolicy(env_name=env.name, **(policy_kwargs or {}))
        self.num_starts = num_starts
        self.num_augment = num_augment
        self.augment = (StateAugmentation(num_augment=num_augment, augment_fn=augment_fn,
                                          first_aug_identity=first_aug_identity, feats=feats)
                        if num_augment > 1 else None)
        for phase in ("train", "val", "test"):          # pomo/model.py:82-85: force multistart decode types
            attr = f"{phase}_decode_type"
            decode_type = getattr(self.policy, attr, None)
            if decode_type is not None and "multistart" not in decode_type:
                setattr(self.policy, attr, f"multistart_{decode_type}")

    def shared_step(self, batch, batch_idx: int = 0, phase: str = "train", dataloader_idx=None, optimizer=None,
                    max_grad_norm: Optional[float] = 1.0) -> dict:
        td = self.env.reset(batch)
        instances = int(td.batch_size[0])
        n_start = self.env.get_num_starts(td) if self.num_starts is None else int(self.num_starts)
        n_aug = 0 if phase == "train" else int(self.num_augment)

        if phase == "train":
            assert n_start > 1, "num_starts must be > 1 during training"
            from .training import reinforce_step
            step = reinforce_step(self.policy, self.env, td, optimizer, baseline="shared",
                                  decode_type=self.policy.train_decode_type, num_starts=n_start,
                                  max_grad_norm=max_grad_norm)
            out = {"reward": step["reward"], "log_likelihood": step["log_likelihood"], "actions": step["actions"],
                   "loss": step["loss"]}
            best = best_of_rollouts(out["reward"], None, instances, 0, n_start)
            out["max_reward"] = best["max_reward"][:, 0]
            return out

        if n_aug > 1:
            td = self.augment(td)
        with torch.inference_mode():
            out = self.policy(td, self.env, phase=phase, num_starts=n_start, return_actions=True)
            if n_start > 1 or n_aug > 1:
                best = best_of_rollouts(out["reward"], out.get("actions"), instances, n_aug, n_start)
        A, S = max(n_aug, 1), max(n_start, 1)
        T = out["actions"].shape[-1]
        # rollout (s*A + a)*B + b  ->  [B, A, S, T] as the reference's unbatchify(actions, (n_aug, n_start)) (a view)
        per_instance = out["actions"].view(S, A, instances, T).permute(2, 1, 0, 3)
        if n_start > 1:
            out["max_reward"] = best["max_reward"] if n_aug > 1 else best["max_reward"][:, 0]
            out["best_multistart_actions"] = (best["best_multistart_actions"] if n_aug > 1
                                              else best["best_multistart_actions"][:, 0])
            out["actions"] = per_instance if n_aug > 1 else per_instance[:, 0]
        if n_aug > 1:
            out["max_aug_reward"] = best["max_aug_reward"]
            out["best_aug_actions"] = best["best_aug_actions"]
        return out
