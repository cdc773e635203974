"""Reward / baseline reductions of the rollout path and their data-parallel form.

Reference: rl4co/models/rl/reinforce/baselines.py:57-61 (SharedBaseline), :162-258
(RolloutBaseline.rollout / mean / t-test), rl4co/models/rl/reinforce/reinforce.py:103-106
(loss), rl4co/models/zoo/pomo/model.py:113-141 (max over starts), and
rl4co/models/rl/common/base.py:224-240 (sync_dist metric means).

Instances are sharded across ranks; all starts of an instance stay on one
rank, so the POMO shared baseline is rank-local exactly as in the reference.  What crosses
ranks is a handful of float64 partial sums produced by rl4co_reward_stats /
rl4co_shared_baseline: one all-reduce (NCCL on GPU ranks, gloo in the CPU tests) per batch.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import _lib


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def all_reduce_sum_(t: torch.Tensor, group=None) -> torch.Tensor:
    """In-place sum over ranks; no-op when not distributed."""
    if _world() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def shard_bounds(total: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of ``total`` instances for ``rank`` (remainder to the first
    ranks)."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reward_stats(reward: torch.Tensor, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Local partials [sum, sum of squares, count] (float64) of a reward vector -
    rl4co_reward_stats.  Accumulates into ``stats`` if given."""
    if not reward.is_cuda:
        raise RuntimeError("reward_stats needs a CUDA tensor (no CPU fallback)")
    if stats is None:
        stats = torch.zeros(3, dtype=torch.float64, device=reward.device)
    r = reward.reshape(-1).contiguous()
    _lib.call("rl4co_reward_stats", r.numel(), _lib.ptr(r, torch.float32), _lib.ptr(stats),
              _lib.stream_ptr(reward.device))
    return stats


def finalize_stats(stats: torch.Tensor, group=None):
    """All-reduce the partials and return (mean, unbiased variance, count) as a float64
    tensor of 3 - host-side logic shared by the GPU path and the gloo tests."""
    stats = all_reduce_sum_(stats.clone(), group)
    n = stats[2]
    mean = stats[0] / n
    var = (stats[1] - n * mean * mean) / torch.clamp(n - 1, min=1)
    return torch.stack((mean, var, n))


def global_reward_mean(reward: torch.Tensor, group=None) -> torch.Tensor:
    """reward.mean() over every rank's shard (RolloutBaseline.mean, baselines.py:186-189;
    the logged val/reward, rl/common/base.py:224-240)."""
    return finalize_stats(reward_stats(reward), group)[0]


def paired_ttest_partials(candidate: torch.Tensor, baseline: torch.Tensor) -> torch.Tensor:
    """Partials of the one-sided paired t-test of RolloutBaseline.epoch_callback
    (baselines.py:202-224): [sum d, sum d^2, n] with d = candidate - baseline."""
    return reward_stats(candidate - baseline)


def paired_ttest_finalize(partials: torch.Tensor, group=None):
    """-> (t statistic, n).  scipy.stats.ttest_rel's t = mean(d) / (std(d, ddof=1)/sqrt(n));
    the p-value lookup stays on the host (baselines.py:215-219)."""
    mean, var, n = finalize_stats(partials, group).tolist()
    t = mean / ((var / n) ** 0.5) if var > 0 else float("nan")
    return t, int(n)


class SharedBaseline:
    """baselines.py:57-61: mean over the starts of each instance + REINFORCE loss partials.
    ``eval(td, reward, ll)`` takes reward / ll in the policy's flat start-major layout
    [S*B] (or [B,S] like the reference after unbatchify) and returns
    (baseline [B,1], loss_partial float64[2] = [sum adv*ll, count])."""

    def __init__(self, num_starts: Optional[int] = None):
        self.num_starts = num_starts

    def eval(self, td, reward: torch.Tensor, ll: Optional[torch.Tensor] = None, env=None,
             num_starts: Optional[int] = None):
        if not reward.is_cuda:
            raise RuntimeError("SharedBaseline.eval needs CUDA tensors (no CPU fallback)")
        dev = reward.device
        if reward.dim() == 2:  # [B,S] -> start-major flat
            B, S = reward.shape
            reward = reward.t().contiguous().view(-1)
            ll = ll.t().contiguous().view(-1) if ll is not None else None
        else:
            S = int(num_starts or self.num_starts)
            B = reward.numel() // S
        out = {"baseline": torch.empty(B, dtype=torch.float32, device=dev),
               "advantage": torch.empty(S * B, dtype=torch.float32, device=dev),
               "max_reward": torch.empty(B, dtype=torch.float32, device=dev),
               "best_start": torch.empty(B, dtype=torch.int64, device=dev),
               "loss_partial": torch.zeros(2, dtype=torch.float64, device=dev)}
        _lib.call("rl4co_shared_baseline", B, S, _lib.ptr(reward.contiguous(), torch.float32),
                  _lib.ptr(ll.contiguous(), torch.float32) if ll is not None else None,
                  _lib.ptr(out["baseline"]), _lib.ptr(out["advantage"]), _lib.ptr(out["max_reward"]),
                  _lib.ptr(out["best_start"]), _lib.ptr(out["loss_partial"]), _lib.stream_ptr(dev))
        self.last = out
        return out["baseline"].unsqueeze(1), out["loss_partial"]


def reinforce_loss(loss_partial: torch.Tensor, group=None) -> torch.Tensor:
    """reinforce.py:103-106 over the GLOBAL batch: -(sum adv*ll)/count after one all-reduce."""
    p = all_reduce_sum_(loss_partial.clone(), group)
    return -(p[0] / p[1])


class RolloutBaseline:
    """baselines.py:162-258, the greedy-rollout consumer of the path: ``rollout`` greedy-rolls
    a dataset shard in batches and ``mean`` is the global mean via one all-reduce."""

    def __init__(self, policy, env, batch_size: int = 65536):
        self.policy, self.env, self.batch_size = policy, env, batch_size
        self.mean = None

    @torch.inference_mode()
    def rollout(self, td_dataset) -> torch.Tensor:
        """baselines.py:226-243: rewards [D_local] of greedy rollouts."""
        D = td_dataset.batch_size[0]
        out = []
        for lo in range(0, D, self.batch_size):
            td = self.env.reset(td_dataset[lo:min(D, lo + self.batch_size)])
            out.append(self.policy(td, self.env, decode_type="greedy")["reward"])
        return torch.cat(out)

    def setup(self, td_dataset, group=None):
        rewards = self.rollout(td_dataset)
        self.mean = global_reward_mean(rewards, group)
        return rewards
