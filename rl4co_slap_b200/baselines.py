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
    """baselines.py:57-61: the mean over the starts of each instance.

    ``eval(td, reward, env=None, on_dim=1) -> (baseline, 0)`` is the reference's surface: ``reward`` is [B,S]
    (after unbatchify, as POMO passes it) and the result is ``reward.mean(on_dim, keepdims=True)``; a flat start-major
    [S*B] reward (what this repo's policy returns) is accepted when ``num_starts`` was given to the constructor and
    yields a flat [S*B] baseline.  ``eval_fused`` is the one-launch variant that also produces the advantage,
    the REINFORCE loss partials and the max over starts (rl4co_shared_baseline)."""

    def __init__(self, num_starts: Optional[int] = None, **kw):
        self.num_starts = num_starts
        self.last = None

    def wrap_dataset(self, dataset, *args, **kw):
        return dataset

    def setup(self, *args, **kw):
        pass

    def epoch_callback(self, *args, **kw):
        pass

    def eval(self, td, reward: torch.Tensor, env=None, on_dim: int = 1):
        if reward.dim() >= 2:
            return reward.mean(dim=on_dim, keepdims=True), 0
        if not self.num_starts:
            raise ValueError("SharedBaseline.eval: a flat [S*B] reward needs SharedBaseline(num_starts=S); "
                             "pass reward as [B,S] otherwise (baselines.py:60-61)")
        baseline, _ = self.eval_fused(reward, None, int(self.num_starts))
        S = int(self.num_starts)
        return baseline.view(1, -1).expand(S, -1).reshape(-1), 0

    def eval_fused(self, reward: torch.Tensor, ll: Optional[torch.Tensor] = None, num_starts: Optional[int] = None):
        """reward / ll flat start-major [S*B] or [B,S] -> (baseline [B,1], loss_partial float64[2] = [sum adv*ll, count]);
        advantage / max_reward / best_start stay in ``self.last``."""
        if not reward.is_cuda:
            raise RuntimeError("SharedBaseline.eval_fused needs CUDA tensors (no CPU fallback)")
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


class NoBaseline:
    """baselines.py:49-53."""

    def __init__(self, *args, **kw):
        pass

    def wrap_dataset(self, dataset, *args, **kw):
        return dataset

    def setup(self, *args, **kw):
        pass

    def epoch_callback(self, *args, **kw):
        pass

    def eval(self, td, reward, env=None):
        return 0, 0


class ExponentialBaseline(NoBaseline):
    """baselines.py:64-83: exponential moving average of the batch-mean reward.  The mean is taken over the GLOBAL
    batch (fp64 partial sums from rl4co_reward_stats, one all-reduce across ranks) - the reference
    averages per rank."""

    def __init__(self, beta: float = 0.8, group=None, **kw):
        self.beta, self.v, self.group = beta, None, group

    def eval(self, td, reward, env=None):
        mean = global_reward_mean(reward, self.group).to(torch.float32)
        v = mean if self.v is None else self.beta * self.v + (1.0 - self.beta) * mean
        self.v = v.detach()
        return self.v, 0


class MeanBaseline:
    """baselines.py:86-90: the batch mean = an exponential baseline with beta 0."""

    def __new__(cls, **kw):
        return ExponentialBaseline(beta=0.0, **kw)


class WarmupBaseline(NoBaseline):
    """baselines.py:93-135: convex combination of ``baseline`` and an exponential baseline, alpha = (epoch + 1) /
    n_epochs during the warm-up epochs."""

    def __init__(self, baseline, n_epochs: int = 1, warmup_exp_beta: float = 0.8, **kw):
        assert n_epochs > 0, "n_epochs to warmup must be positive"
        self.baseline = baseline
        self.warmup_baseline = ExponentialBaseline(warmup_exp_beta)
        self.alpha, self.n_epochs = 0, n_epochs

    def wrap_dataset(self, dataset, *args, **kw):
        if self.alpha > 0:
            return self.baseline.wrap_dataset(dataset, *args, **kw)
        return self.warmup_baseline.wrap_dataset(dataset, *args, **kw)

    def setup(self, *args, **kw):
        self.baseline.setup(*args, **kw)

    def eval(self, td, reward, env=None):
        if self.alpha == 1:
            return self.baseline.eval(td, reward, env)
        if self.alpha == 0:
            return self.warmup_baseline.eval(td, reward, env)
        v_b, l_b = self.baseline.eval(td, reward, env)
        v_wb, l_wb = self.warmup_baseline.eval(td, reward, env)
        return self.alpha * v_b + (1 - self.alpha) * v_wb, self.alpha * l_b + (1 - self.alpha) * l_wb

    def epoch_callback(self, *args, **kw):
        self.baseline.epoch_callback(*args, **kw)
        if kw["epoch"] < self.n_epochs:
            self.alpha = (kw["epoch"] + 1) / float(self.n_epochs)


def reinforce_loss(loss_partial: torch.Tensor, group=None) -> torch.Tensor:
    """reinforce.py:103-106 over the GLOBAL batch: -(sum adv*ll)/count after one all-reduce."""
    p = all_reduce_sum_(loss_partial.clone(), group)
    return -(p[0] / p[1])


class RolloutBaseline:
    """baselines.py:162-258, the greedy-rollout consumer of the path, on device-resident datasets.

    Same surface as the reference (``setup`` / ``_update_policy`` / ``eval`` / ``epoch_callback`` /
    ``rollout`` / ``wrap_dataset``); differences: the evaluation dataset is one device TensorDict sliced
    in place (no DataLoader / host collate), rewards stay on the device, and with ``group`` (or an
    initialised default process group) every rank rolls out its own contiguous shard of the dataset while
    the mean and the paired t-test come from one all-reduce of three fp64 partial sums each."""

    def __init__(self, bl_alpha: float = 0.05, **kw):
        self.bl_alpha = bl_alpha
        self.policy = self.dataset = self.bl_vals = self.mean = None
        self.group = kw.get("group")

    def setup(self, *args, **kw):
        self._update_policy(*args, **kw)

    def _update_policy(self, policy, env, batch_size: int = 65536, device="cuda", dataset_size=None, dataset=None):
        """baselines.py:176-189: freeze a copy of the policy and greedy-roll it over the evaluation dataset."""
        import copy
        self.policy = copy.deepcopy(policy).to(device)
        if dataset is None:
            self.dataset = env.dataset(batch_size=[dataset_size])
        else:       # (the reference ignores a passed dataset here and keeps the old one; accept it)
            self.dataset = dataset
        self.bl_vals = self.rollout(self.policy, env, batch_size, device, self.dataset)
        self.mean = global_reward_mean(self.bl_vals, self.group).item()

    def eval(self, td, reward, env):
        """baselines.py:191-200: greedy rollout of the frozen policy on the batch."""
        with torch.inference_mode():
            reward = self.policy(td, env)["reward"]
        return reward, 0

    def epoch_callback(self, policy, env, batch_size: int = 65536, device="cuda", epoch=None, dataset_size=None):
        """baselines.py:202-224: challenge the baseline; replace it if the candidate is better with
        p < bl_alpha in a one-sided paired t-test.  Returns (candidate_mean, t, p_val, updated)."""
        candidate_vals = self.rollout(policy, env, batch_size, device)
        candidate_mean = global_reward_mean(candidate_vals, self.group).item()
        t = p_val = None
        updated = False
        if candidate_mean - self.mean > 0:
            # ttest_rel(-candidate, -baseline): d = baseline - candidate, t must be negative
            t, n = paired_ttest_finalize(paired_ttest_partials(self.bl_vals, candidate_vals), self.group)
            from scipy.stats import t as student_t
            p_val = float(student_t.sf(abs(t), n - 1))          # = two-sided p / 2
            assert t < 0, "T-statistic should be negative"
            if p_val < self.bl_alpha:
                self._update_policy(policy, env, batch_size, device, dataset_size)
                updated = True
        return candidate_mean, t, p_val, updated

    def _world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def _rank(self) -> int:
        return dist.get_rank(self.group) if dist.is_available() and dist.is_initialized() else 0

    def rollout(self, policy, env, batch_size: int = 65536, device="cuda", dataset=None, shard: bool = True) -> torch.Tensor:
        """baselines.py:226-243: rewards of greedy rollouts, in dataset order, on the device.  With a process group
        every rank holds the SAME dataset (same seed) and rolls out only its contiguous shard -> [D_local]."""
        from .data import TensorDictDataset
        dataset = self.dataset if dataset is None else dataset
        if not isinstance(dataset, TensorDictDataset):
            dataset = TensorDictDataset(dataset)
        if shard and self._world() > 1:      # this rank's contiguous slice; the statistics are all-reduced
            lo, hi = shard_bounds(len(dataset), self._rank(), self._world())
            dataset = TensorDictDataset(dataset.data[lo:hi])
        policy.eval()
        policy = policy.to(device)
        out = []
        with torch.inference_mode():
            for batch in dataset.batches(batch_size):
                td = env.reset(batch.to(device))
                out.append(policy(td, env, decode_type="greedy")["reward"])
        return torch.cat(out, 0)

    def wrap_dataset(self, dataset, env, batch_size: int = 65536, device="cuda", **kw):
        """baselines.py:245-258: attach the frozen policy's greedy rewards as ``"extra"``."""
        from .data import TensorDictDataset
        if not isinstance(dataset, TensorDictDataset):
            dataset = TensorDictDataset(dataset)
        # the wrapped (training) dataset is this rank's own data: every row needs its baseline value here
        rewards = self.rollout(self.policy, env, batch_size, device, dataset=dataset, shard=False).detach()
        return dataset.add_key("extra", rewards)


# baselines.py:275-299 (the critic baseline needs a critic network: out of scope)
REINFORCE_BASELINES_REGISTRY = {"no": NoBaseline, "shared": SharedBaseline, "exponential": ExponentialBaseline,
                                "mean": MeanBaseline, "rollout_only": RolloutBaseline, "warmup": WarmupBaseline}


def get_reinforce_baseline(name, **kw):
    """baselines.py:286-299: "rollout" = a warm-up baseline around the greedy-rollout baseline."""
    if name == "warmup":
        inner = kw.get("baseline", "rollout")
        if not isinstance(inner, NoBaseline) and isinstance(inner, str):
            inner = get_reinforce_baseline(inner, **kw)
        return WarmupBaseline(inner, **{k: v for k, v in kw.items() if k != "baseline"})
    if name == "rollout":
        return WarmupBaseline(RolloutBaseline(bl_alpha=kw.get("bl_alpha", 0.05), group=kw.get("group")),
                              kw.get("n_epochs", 1), kw.get("exp_beta", 0.8))
    if name is None:
        name = "no"
    cls = REINFORCE_BASELINES_REGISTRY.get(name)
    if cls is None:
        raise ValueError(f"Unknown baseline {name}. Available baselines: {REINFORCE_BASELINES_REGISTRY.keys()}")
    return cls(**kw)
