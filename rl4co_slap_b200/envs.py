"""TSP / CVRP / SLAP environments with the reference's API over the CUDA kernels.

Mirrors RL4COEnvBase (rl4co/envs/common/base.py:19-213): ``reset(td|batch_size)``,
``step(td) -> {"next": td}``, ``_step``, ``get_action_mask``, ``get_reward(td, actions)``,
``check_solution_validity``, ``get_num_starts``, ``select_start_nodes``, ``generator``.
State tensors keep the reference's keys, dtypes and shapes so a td produced
by either implementation is consumable by the other.

Every transition / mask / reward / validity computation is a launch into the C-ABI library
(include/rl4co_slap_b200.h).  There is no CPU path: tensors must live on a CUDA device.
``replay(td, actions)`` is the fused form of ``reset`` + T x ``_step`` + ``get_reward``.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib
from .generators import ATSPGenerator, CVRPGenerator, OPGenerator, PCTSPGenerator, SLAPGenerator, TSPGenerator
from .ops import get_num_starts, select_start_nodes
from .tensordict_compat import TensorDict

_u8 = torch.uint8


def _norm2(d: torch.Tensor) -> torch.Tensor:
    """L2 norm of [..., 2] differences with the rounding sequence of torch's CPU vector norm,
    sqrt(fma(dy, dy, dx * dx)) (addcmul is a fused multiply-add on CUDA): the reset-time budgets the masks
    are compared against are then the same fp32 numbers as the CPU reference's."""
    dx, dy = d[..., 0], d[..., 1]
    return torch.sqrt(torch.addcmul(dx * dx, dy, dy))


def _as_u8(t: torch.Tensor) -> torch.Tensor:
    """bool tensors are one byte per element: reinterpret, no copy."""
    return t.view(_u8) if t.dtype == torch.bool else t


def _need_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"{what}: rl4co_slap_b200 environments run on CUDA tensors only (no CPU fallback); "
            f"got device {t.device}")


class RL4COEnvBase:
    """envs/common/base.py:19-213 (the hot-path surface; dataset / torchrl-spec plumbing is
    out of scope)."""

    name = "base"
    batch_locked = False

    def __init__(self, *, check_solution: bool = True, seed: Optional[int] = None,
                 device: str = "cuda", batch_size=None, **kwargs):
        kwargs.pop("name", None)
        self.unused_kwargs = dict(kwargs)  # the reference only logs these (base.py:72-78)
        self.check_solution = check_solution
        self.device = torch.device(device) if device is not None else None
        self.batch_size = torch.Size([]) if batch_size is None else torch.Size(batch_size)
        if seed is None:
            seed = torch.empty((), dtype=torch.int64).random_().item()
        self.set_seed(seed)

    # -- torchrl EnvBase surface used by the path ------------------------------------
    def set_seed(self, seed: Optional[int]) -> None:
        self._seed = seed
        self.rng = torch.manual_seed(seed)

    def to(self, device):
        self.device = torch.device(device)
        return self

    def step(self, td):
        """base.py:121-133 (non-torchrl mode)."""
        return {"next": self._step(td)}

    def reset(self, td=None, batch_size=None):
        """base.py:135-143 + torchrl EnvBase.reset: generate if empty, call ``_reset``, add
        done / terminated [B,1] False, merge the result into the input td."""
        if batch_size is None:
            batch_size = self.batch_size if td is None else td.batch_size
        if td is None or td.is_empty():
            td = self.generator(batch_size=batch_size)
            if self.device is not None and self.device.type == "cuda" and td.device != self.device:
                td = td.to(self.device)
        batch_size = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        self.to(td.device)
        out = self._reset(td, batch_size=batch_size)
        dev = td.device
        for k in ("done", "terminated"):
            if k not in out:
                out.set(k, torch.zeros((*batch_size, 1), dtype=torch.bool, device=dev))
        merged = td.clone(False) if hasattr(td, "clone") else td
        merged.update(out)
        return merged

    def get_reward(self, td, actions):
        """base.py:182-188."""
        if self.check_solution:
            _lib.raise_deferred(actions.device)  # out-of-range actions seen by earlier _step calls
            self.check_solution_validity(td, actions)
        return self._get_reward(td, actions)

    def dataset(self, batch_size=[], phase: str = "train", filename: Optional[str] = None):
        """base.py:236-273: a dataset of observations, loaded from ``filename`` / ``{phase}_file`` when one is
        set (missing file -> generated, as in the reference), else generated - here straight on the env's device."""
        from .data import TensorDictDataset, generate_dataset
        f = getattr(self, f"{phase}_file", None) if filename is None else filename
        if isinstance(batch_size, (list, tuple, torch.Size)):
            n = int(batch_size[0]) if len(batch_size) else None
        else:
            n = int(batch_size)
        td = None
        if f is not None:
            try:
                td = self.load_data(f, batch_size)
                if self.device is not None and self.device.type == "cuda":
                    from .data import to_device
                    td = to_device(td, self.device)
            except FileNotFoundError:
                td = None
        if td is None:
            if n is None:
                raise ValueError("dataset(): batch_size=[n] is needed to generate a dataset")
            td = generate_dataset(self, n)
        return TensorDictDataset(td)

    def load_data(self, fpath, batch_size=[]):
        """base.py:270-273 / cvrp/env.py:192-199."""
        from .data import load_data
        return load_data(self.name, fpath)

    def get_num_starts(self, td):
        return get_num_starts(td, self.name)

    def select_start_nodes(self, td, num_starts):
        return select_start_nodes(td, self, num_starts)

    def get_action_mask(self, td):
        raise NotImplementedError

    def check_solution_validity(self, td, actions):
        raise NotImplementedError

    def _step(self, td):
        raise NotImplementedError

    def _reset(self, td, batch_size=None):
        raise NotImplementedError

    def _get_reward(self, td, actions):
        raise NotImplementedError


# =====================================================================================
class TSPEnv(RL4COEnvBase):
    """rl4co/envs/routing/tsp/env.py:29-201."""

    name = "tsp"

    def __init__(self, generator: TSPGenerator = None, generator_params: dict = {}, **kwargs):
        super().__init__(**kwargs)
        if generator is None:
            generator = TSPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator

    def _reset(self, td, batch_size=None):
        """tsp/env.py:95-120 -> rl4co_tsp_reset."""
        locs = td["locs"]
        _need_cuda(locs, "TSPEnv.reset")
        dev, B, N = locs.device, int(batch_size[0]), locs.shape[-2]
        cur = torch.empty(B, dtype=torch.int64, device=dev)
        i = torch.empty(B, 1, dtype=torch.int64, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        reward = torch.empty(B, 1, dtype=torch.float32, device=dev)
        done = torch.empty(B, 1, dtype=torch.bool, device=dev)
        term = torch.empty(B, 1, dtype=torch.bool, device=dev)
        _lib.call("rl4co_tsp_reset", B, N, None, _lib.ptr(cur), _lib.ptr(i), _lib.ptr(_as_u8(mask)),
                  _lib.ptr(reward), _lib.ptr(_as_u8(done)), _lib.ptr(_as_u8(term)), _lib.stream_ptr(dev))
        # first_node and current_node are the SAME tensor object in the reference (:103,111-112)
        return TensorDict({"locs": locs, "first_node": cur, "current_node": cur, "i": i,
                           "action_mask": mask, "reward": reward, "done": done, "terminated": term},
                          batch_size=batch_size)

    @staticmethod
    def _step(td):
        """tsp/env.py:67-93 -> rl4co_tsp_step.  Out of place on the mask like the reference's
        scatter; `done` and `reward` come back as [B] (reward = zeros_like(done), bool)."""
        action = td["action"]
        _need_cuda(action, "TSPEnv._step")
        dev, B = action.device, action.shape[0]
        mask = td["action_mask"].clone()
        N = mask.shape[-1]
        i = td["i"].clone()
        first = td["first_node"].clone()
        cur = torch.empty_like(action)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        flag = torch.empty(1, dtype=torch.int32, device=dev)
        status = _lib.deferred_status(dev)
        _lib.call("rl4co_tsp_step", B, N, _lib.ptr(action.contiguous(), torch.int64), _lib.ptr(i),
                  _lib.ptr(first), _lib.ptr(cur), _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)),
                  _lib.ptr(flag), _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"first_node": first, "current_node": cur, "i": i, "action_mask": mask,
                   "reward": torch.zeros_like(done), "done": done})
        return td

    def get_action_mask(self, td):
        return td["action_mask"]

    def replay(self, td, actions, check: Optional[bool] = None):
        """Fused reset + T x _step + get_reward (rl4co_tsp_replay): final state + reward in
        one launch.  Returns (td_final, reward [B])."""
        locs = td["locs"]
        _need_cuda(locs, "TSPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        T = actions.shape[1]
        check = self.check_solution if check is None else check
        first = torch.empty(B, dtype=torch.int64, device=dev)
        cur = torch.empty(B, dtype=torch.int64, device=dev)
        i = torch.empty(B, 1, dtype=torch.int64, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        _lib.call("rl4co_tsp_replay", B, N, T, _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), _lib.ptr(first), _lib.ptr(cur), _lib.ptr(i),
                  _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)), _lib.ptr(reward), int(bool(check)),
                  _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)  # one D2H read: validity bits (if checked) + action range
        out = TensorDict({"locs": locs, "first_node": first, "current_node": cur, "i": i,
                          "action_mask": mask, "reward": torch.zeros_like(done), "done": done},
                         batch_size=[B])
        return out, reward

    def _get_reward(self, td, actions):
        """tsp/env.py:157-163 (note: it re-checks validity itself when check_solution)."""
        return self._reward_and_check(td, actions, self.check_solution)

    def _reward_and_check(self, td, actions, check):
        locs = td["locs"]
        _need_cuda(locs, "TSPEnv.get_reward")
        dev, (B, N, _) = locs.device, locs.shape
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        _lib.call("rl4co_tsp_replay", B, N, actions.shape[1], _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), None, None, None, None, None,
                  _lib.ptr(reward), int(bool(check)), _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)  # one D2H read: validity bits (if checked) + action range
        return reward

    def check_solution_validity(self, td, actions) -> None:
        """tsp/env.py:165-173: AssertionError("Invalid tour")."""
        self._reward_and_check(td, actions, True)


# =====================================================================================
class ATSPEnv(TSPEnv):
    """rl4co/envs/routing/atsp/env.py:22-175: the TSP state machine on a cost matrix.  Environment only - the
    reference registers TSPInitEmbedding for "atsp", which needs coordinates the instances do not have."""

    name = "atsp"

    def __init__(self, generator: ATSPGenerator = None, generator_params: dict = {}, **kwargs):
        RL4COEnvBase.__init__(self, **kwargs)
        if generator is None:
            generator = ATSPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator

    def _reset(self, td, batch_size=None):
        """atsp/env.py:92-112 (current_node / first_node are ONE [B,1] tensor there)."""
        cm = td["cost_matrix"]
        _need_cuda(cm, "ATSPEnv.reset")
        dev, B, N = cm.device, int(batch_size[0]), cm.shape[-1]
        cur = torch.zeros(B, 1, dtype=torch.int64, device=dev)
        return TensorDict({"cost_matrix": cm, "first_node": cur, "current_node": cur,
                           "i": torch.zeros(B, 1, dtype=torch.int64, device=dev),
                           "action_mask": torch.ones(B, N, dtype=torch.bool, device=dev)}, batch_size=batch_size)

    @staticmethod
    def _step(td):
        """atsp/env.py:64-90 -> rl4co_tsp_step (same transition; first_node becomes [B] like the action)."""
        td.set("first_node", td["first_node"].reshape(-1))
        return TSPEnv._step(td)

    def _replay_call(self, td, actions, check, want_state):
        cm = td["cost_matrix"]
        _need_cuda(cm, "ATSPEnv.replay")
        dev, (B, N, _) = cm.device, cm.shape
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        st = {}
        if want_state:
            st = {"first_node": torch.empty(B, dtype=torch.int64, device=dev),
                  "current_node": torch.empty(B, dtype=torch.int64, device=dev),
                  "i": torch.empty(B, 1, dtype=torch.int64, device=dev),
                  "action_mask": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "done": torch.empty(B, dtype=torch.bool, device=dev)}
        p = lambda k: _lib.ptr(_as_u8(st[k])) if k in st else None  # noqa: E731
        _lib.call("rl4co_atsp_replay", B, N, actions.shape[1], _lib.ptr(cm.contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), p("first_node"), p("current_node"), p("i"),
                  p("action_mask"), p("done"), _lib.ptr(reward), int(bool(check)), _lib.ptr(status),
                  _lib.stream_ptr(dev))
        _lib.raise_for_status(status)
        return st, reward

    def replay(self, td, actions, check: Optional[bool] = None):
        check = self.check_solution if check is None else check
        st, reward = self._replay_call(td, actions, check, True)
        out = TensorDict({"cost_matrix": td["cost_matrix"], **st, "reward": torch.zeros_like(st["done"])},
                         batch_size=[actions.shape[0]])
        return out, reward

    def _get_reward(self, td, actions):
        """atsp/env.py:152-162."""
        return self._replay_call(td, actions, False, False)[1]

    def check_solution_validity(self, td, actions) -> None:
        """atsp/env.py:164-171: AssertionError("Invalid tour")."""
        self._replay_call(td, actions, True, False)


# =====================================================================================
class CVRPEnv(RL4COEnvBase):
    """rl4co/envs/routing/cvrp/env.py:29-258."""

    name = "cvrp"

    def __init__(self, generator: CVRPGenerator = None, generator_params: dict = {}, **kwargs):
        super().__init__(**kwargs)
        if generator is None:
            generator = CVRPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator

    def _reset(self, td, batch_size=None):
        """cvrp/env.py:107-135."""
        _need_cuda(td["locs"], "CVRPEnv.reset")
        dev, B = td["locs"].device, int(batch_size[0])
        locs = torch.cat((td["depot"][:, None, :], td["locs"]), -2)
        N = locs.shape[-2]
        out = TensorDict({
            "locs": locs, "demand": td["demand"],
            "current_node": torch.zeros(B, 1, dtype=torch.long, device=dev),
            "used_capacity": torch.zeros(B, 1, device=dev),
            "vehicle_capacity": torch.full((B, 1), self.generator.vehicle_capacity, device=dev),
            "visited": torch.zeros(B, N, dtype=torch.uint8, device=dev),
        }, batch_size=batch_size)
        out.set("action_mask", self.get_action_mask(out))
        return out

    @staticmethod
    def get_action_mask(td):
        """cvrp/env.py:137-149 -> rl4co_cvrp_action_mask."""
        demand = td["demand"]
        _need_cuda(demand, "CVRPEnv.get_action_mask")
        dev, (B, n) = demand.device, demand.shape
        mask = torch.empty(B, n + 1, dtype=torch.bool, device=dev)
        _lib.call("rl4co_cvrp_action_mask", B, n + 1, _lib.ptr(demand.contiguous(), torch.float32),
                  _lib.ptr(td["used_capacity"].contiguous(), torch.float32),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32),
                  _lib.ptr(td["visited"].contiguous(), _u8),
                  _lib.ptr(td["current_node"].contiguous(), torch.int64),
                  _lib.ptr(_as_u8(mask)), _lib.stream_ptr(dev))
        return mask

    def _step(self, td):
        """cvrp/env.py:73-105 -> rl4co_cvrp_step (transition + mask in one launch)."""
        action = td["action"]
        _need_cuda(action, "CVRPEnv._step")
        dev, B = action.device, action.shape[0]
        demand = td["demand"]
        N = demand.shape[-1] + 1
        used = td["used_capacity"].clone()
        visited = td["visited"].clone()
        cur = torch.empty(B, 1, dtype=torch.int64, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        status = _lib.deferred_status(dev)
        _lib.call("rl4co_cvrp_step", B, N, _lib.ptr(action.contiguous(), torch.int64),
                  _lib.ptr(demand.contiguous(), torch.float32),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32), _lib.ptr(used),
                  _lib.ptr(cur), _lib.ptr(visited), _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)),
                  _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"current_node": cur, "used_capacity": used, "visited": visited,
                   "reward": torch.zeros_like(done), "done": done})
        td.set("action_mask", mask)
        return td

    def _replay_call(self, td, actions, check, want_state):
        locs, demand = td["locs"], td["demand"]
        _need_cuda(locs, "CVRPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        T = actions.shape[1]
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        st = {}
        if want_state:
            st = {"current_node": torch.empty(B, 1, dtype=torch.int64, device=dev),
                  "used_capacity": torch.empty(B, 1, dtype=torch.float32, device=dev),
                  "visited": torch.empty(B, N, dtype=_u8, device=dev),
                  "action_mask": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "done": torch.empty(B, dtype=torch.bool, device=dev)}
        p = lambda k: _lib.ptr(_as_u8(st[k])) if k in st else None  # noqa: E731
        _lib.call("rl4co_cvrp_replay", B, N, T, _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(demand.contiguous(), torch.float32),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), p("current_node"), p("used_capacity"),
                  p("visited"), p("action_mask"), p("done"), _lib.ptr(reward), int(bool(check)),
                  _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)  # one D2H read: validity bits (if checked) + action range
        return st, reward

    def replay(self, td, actions, check: Optional[bool] = None):
        """Fused reset + T x _step + get_reward.  ``td`` is a reset td (locs incl. depot)."""
        check = self.check_solution if check is None else check
        st, reward = self._replay_call(td, actions, check, True)
        out = TensorDict({"locs": td["locs"], "demand": td["demand"],
                          "vehicle_capacity": td["vehicle_capacity"], **st,
                          "reward": torch.zeros_like(st["done"])}, batch_size=[actions.shape[0]])
        return out, reward

    def _get_reward(self, td, actions):
        """cvrp/env.py:151-160."""
        return self._replay_call(td, actions, False, False)[1]

    def check_solution_validity(self, td, actions) -> None:
        """cvrp/env.py:162-190: "Invalid tour" / "Used more than capacity"."""
        self._replay_call(td, actions, True, False)


# =====================================================================================
class SDVRPEnv(CVRPEnv):
    """rl4co/envs/routing/sdvrp/env.py:22-214 (split-delivery VRP): customers may be visited several times, each
    visit delivers min(remaining demand, remaining capacity).  Generator and reward are the CVRP ones."""

    name = "sdvrp"

    def _reset(self, td, batch_size=None):
        """sdvrp/env.py:103-130."""
        _need_cuda(td["locs"], "SDVRPEnv.reset")
        dev, B = td["locs"].device, int(batch_size[0])
        out = TensorDict({
            "locs": torch.cat((td["depot"][..., None, :], td["locs"]), -2),
            "demand": td["demand"],
            "demand_with_depot": torch.cat((torch.zeros_like(td["demand"][..., 0:1]), td["demand"]), -1),
            "current_node": torch.zeros(B, 1, dtype=torch.long, device=dev),
            "used_capacity": torch.zeros(B, 1, device=dev),
            "vehicle_capacity": torch.full((B, 1), self.generator.vehicle_capacity, device=dev),
        }, batch_size=batch_size)
        out.set("action_mask", self.get_action_mask(out))
        return out

    @staticmethod
    def get_action_mask(td):
        """sdvrp/env.py:132-140 -> rl4co_sdvrp_action_mask."""
        dwd = td["demand_with_depot"]
        _need_cuda(dwd, "SDVRPEnv.get_action_mask")
        dev, (B, N) = dwd.device, dwd.shape
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        _lib.call("rl4co_sdvrp_action_mask", B, N, _lib.ptr(dwd.contiguous(), torch.float32),
                  _lib.ptr(td["used_capacity"].contiguous(), torch.float32),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32),
                  _lib.ptr(td["current_node"].contiguous(), torch.int64), _lib.ptr(_as_u8(mask)), _lib.stream_ptr(dev))
        return mask

    def _step(self, td):
        """sdvrp/env.py:63-101 -> rl4co_sdvrp_step (transition + mask in one launch)."""
        action = td["action"]
        _need_cuda(action, "SDVRPEnv._step")
        dev, (B, N) = action.device, td["demand_with_depot"].shape
        dwd, used = td["demand_with_depot"].clone(), td["used_capacity"].clone()
        cur = torch.empty(B, 1, dtype=torch.int64, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        status = _lib.deferred_status(dev)
        _lib.call("rl4co_sdvrp_step", B, N, _lib.ptr(action.contiguous(), torch.int64),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32), _lib.ptr(dwd), _lib.ptr(used),
                  _lib.ptr(cur), _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)), _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"demand_with_depot": dwd, "current_node": cur, "used_capacity": used,
                   "reward": torch.zeros_like(done), "done": done})
        td.set("action_mask", mask)
        return td

    def _replay_call(self, td, actions, check, want_state):
        locs, demand = td["locs"], td["demand"]
        _need_cuda(locs, "SDVRPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        st = {}
        if want_state:
            st = {"current_node": torch.empty(B, 1, dtype=torch.int64, device=dev),
                  "used_capacity": torch.empty(B, 1, dtype=torch.float32, device=dev),
                  "demand_with_depot": torch.empty(B, N, dtype=torch.float32, device=dev),
                  "action_mask": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "done": torch.empty(B, dtype=torch.bool, device=dev)}
        p = lambda k: _lib.ptr(_as_u8(st[k])) if k in st else None  # noqa: E731
        _lib.call("rl4co_sdvrp_replay", B, N, actions.shape[1], _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(demand.contiguous(), torch.float32),
                  _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), p("current_node"), p("used_capacity"),
                  p("demand_with_depot"), p("action_mask"), p("done"), _lib.ptr(reward), int(bool(check)),
                  _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)
        return st, reward

    def replay(self, td, actions, check: Optional[bool] = None):
        check = self.check_solution if check is None else check
        st, reward = self._replay_call(td, actions, check, True)
        out = TensorDict({"locs": td["locs"], "demand": td["demand"], "vehicle_capacity": td["vehicle_capacity"], **st,
                          "reward": torch.zeros_like(st["done"])}, batch_size=[actions.shape[0]])
        return out, reward

    def check_solution_validity(self, td, actions) -> None:
        """sdvrp/env.py:142-164: "Cannot visit depot twice if any nonzero demand" / "All demand must be satisfied"."""
        self._replay_call(td, actions, True, False)


# =====================================================================================
class OPEnv(RL4COEnvBase):
    """rl4co/envs/routing/op/env.py:25-238 (orienteering problem): collect prizes and be back at the depot
    within ``max_length``; the sibling routing environment on the same kernels' skeleton."""

    name = "op"

    def __init__(self, generator: OPGenerator = None, generator_params: dict = {}, prize_type: str = "dist",
                 **kwargs):
        super().__init__(**kwargs)
        if generator is None:
            generator = OPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator
        assert prize_type in ("dist", "unif", "const"), f"Invalid prize_type: {prize_type}"
        self.prize_type = prize_type

    def _reset(self, td, batch_size=None):
        """op/env.py:110-152.  The per-node budget is plain tensor arithmetic done once per episode."""
        _need_cuda(td["locs"], "OPEnv.reset")
        dev, B = td["locs"].device, int(batch_size[0])
        locs = torch.cat((td["depot"][:, None, :], td["locs"]), -2)
        N = locs.shape[-2]
        out = TensorDict({
            "locs": locs,
            "prize": F.pad(td["prize"], (1, 0), mode="constant", value=0),
            "tour_length": torch.zeros(B, device=dev),
            "max_length": td["max_length"][..., None] - _norm2(td["depot"][..., None, :] - locs) - 1e-6,
            "current_node": torch.zeros(B, 1, dtype=torch.long, device=dev),
            "visited": torch.zeros(B, N, dtype=torch.bool, device=dev),
            "current_total_prize": torch.zeros(B, dtype=torch.float, device=dev),
            "i": torch.zeros(B, dtype=torch.int64, device=dev),
        }, batch_size=batch_size)
        out.set("action_mask", self.get_action_mask(out))
        return out

    @staticmethod
    def get_action_mask(td):
        """op/env.py:154-170 -> rl4co_op_action_mask."""
        locs = td["locs"]
        _need_cuda(locs, "OPEnv.get_action_mask")
        dev, (B, N, _) = locs.device, locs.shape
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        _lib.call("rl4co_op_action_mask", B, N, _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(td["max_length"].contiguous(), torch.float32),
                  _lib.ptr(td["tour_length"].contiguous(), torch.float32),
                  _lib.ptr(_as_u8(td["visited"].contiguous())),
                  _lib.ptr(td["current_node"].contiguous(), torch.int64),
                  _lib.ptr(_as_u8(mask)), _lib.stream_ptr(dev))
        return mask

    def _step(self, td):
        """op/env.py:74-108 -> rl4co_op_step (transition + mask in one launch)."""
        action = td["action"]
        _need_cuda(action, "OPEnv._step")
        dev, (B, N, _) = action.device, td["locs"].shape
        length, total = td["tour_length"].clone(), td["current_total_prize"].clone()
        visited, i = td["visited"].clone(), td["i"].clone()
        cur = td["current_node"].clone()
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        status = _lib.deferred_status(dev)
        _lib.call("rl4co_op_step", B, N, _lib.ptr(action.contiguous(), torch.int64),
                  _lib.ptr(td["locs"].contiguous(), torch.float32), _lib.ptr(td["prize"].contiguous(), torch.float32),
                  _lib.ptr(td["max_length"].contiguous(), torch.float32), _lib.ptr(length), _lib.ptr(total),
                  _lib.ptr(cur, torch.int64), _lib.ptr(_as_u8(visited)), _lib.ptr(i, torch.int64),
                  _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)), _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"tour_length": length, "current_node": cur, "visited": visited,
                   "current_total_prize": total, "i": i, "reward": torch.zeros_like(done), "done": done})
        td.set("action_mask", mask)
        return td

    def _replay_call(self, td, actions, check, want_state):
        locs = td["locs"]
        _need_cuda(locs, "OPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        T = actions.shape[1]
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        st = {}
        if want_state:
            st = {"current_node": torch.empty(B, 1, dtype=torch.int64, device=dev),
                  "tour_length": torch.empty(B, dtype=torch.float32, device=dev),
                  "current_total_prize": torch.empty(B, dtype=torch.float32, device=dev),
                  "visited": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "action_mask": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "i": torch.empty(B, dtype=torch.int64, device=dev),
                  "done": torch.empty(B, dtype=torch.bool, device=dev)}
        p = lambda k: _lib.ptr(_as_u8(st[k])) if k in st else None  # noqa: E731
        _lib.call("rl4co_op_replay", B, N, T, _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(td["prize"].contiguous(), torch.float32),
                  _lib.ptr(td["max_length"].contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), p("current_node"), p("tour_length"),
                  p("current_total_prize"), p("visited"), p("action_mask"), p("i"), p("done"),
                  _lib.ptr(reward), int(bool(check)), _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)
        return st, reward

    def replay(self, td, actions, check: Optional[bool] = None):
        """Fused reset-state + T x _step + get_reward.  ``td`` is a reset td."""
        check = self.check_solution if check is None else check
        st, reward = self._replay_call(td, actions, check, True)
        out = TensorDict({"locs": td["locs"], "prize": td["prize"], "max_length": td["max_length"], **st,
                          "reward": torch.zeros_like(st["done"])}, batch_size=[actions.shape[0]])
        return out, reward

    def _get_reward(self, td, actions):
        """op/env.py:172-182."""
        if actions.size(-1) == 1:
            assert (actions == 0).all(), "If all length 1 tours, they should be zero"
            return torch.zeros(actions.size(0), dtype=torch.float, device=actions.device)
        return self._replay_call(td, actions, False, False)[1]

    def check_solution_validity(self, td, actions, add_distance_to_depot: bool = True) -> None:
        """op/env.py:184-215: "Duplicates" / "Max length exceeded" (``td`` carries the reset state's budgets, so
        the depot leg is added back as with the reference's default add_distance_to_depot=True)."""
        if not add_distance_to_depot:
            raise NotImplementedError("check_solution_validity(add_distance_to_depot=False)")
        self._replay_call(td, actions, True, False)


# =====================================================================================
class PCTSPEnv(RL4COEnvBase):
    """rl4co/envs/routing/pctsp/env.py:26-238 (prize-collecting TSP): collect a total prize of at least 1 (or visit
    everything), pay the penalties of the customers left out."""

    name = "pctsp"
    _stochastic = False

    def __init__(self, generator: PCTSPGenerator = None, generator_params: dict = {}, **kwargs):
        super().__init__(**kwargs)
        if generator is None:
            generator = PCTSPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator

    @property
    def stochastic(self):
        return self._stochastic

    def _reset(self, td, batch_size=None):
        """pctsp/env.py:103-139."""
        _need_cuda(td["locs"], "PCTSPEnv.reset")
        dev, B = td["locs"].device, int(batch_size[0])
        locs = torch.cat([td["depot"][..., None, :], td["locs"]], dim=-2)
        real = td["stochastic_prize"] if self.stochastic else td["deterministic_prize"]
        penalty = td["penalty"]
        out = TensorDict({
            "locs": locs,
            "current_node": torch.zeros(B, dtype=torch.int64, device=dev),
            "expected_prize": td["deterministic_prize"],
            "real_prize": torch.cat([torch.zeros_like(real[..., :1]), real], dim=-1),
            "penalty": F.pad(penalty, (1, 0), mode="constant", value=0),
            "cur_total_prize": torch.zeros(B, device=dev),
            "cur_total_penalty": penalty.sum(-1),
            "visited": torch.zeros(B, locs.shape[-2], dtype=torch.bool, device=dev),
            "prize_required": torch.full((B,), float(self.generator.prize_required), device=dev),
            "i": torch.zeros(B, dtype=torch.int64, device=dev),
        }, batch_size=batch_size)
        out.set("action_mask", self.get_action_mask(out))
        return out

    @staticmethod
    def get_action_mask(td):
        """pctsp/env.py:141-148 -> rl4co_pctsp_action_mask."""
        visited = td["visited"]
        _need_cuda(visited, "PCTSPEnv.get_action_mask")
        dev, (B, N) = visited.device, visited.shape
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        _lib.call("rl4co_pctsp_action_mask", B, N, _lib.ptr(td["cur_total_prize"].contiguous(), torch.float32),
                  _lib.ptr(_as_u8(visited.contiguous())), _lib.ptr(_as_u8(mask)), _lib.stream_ptr(dev))
        return mask

    def _step(self, td):
        """pctsp/env.py:69-101 -> rl4co_pctsp_step (transition + mask in one launch)."""
        action = td["action"]
        _need_cuda(action, "PCTSPEnv._step")
        dev, (B, N) = action.device, td["visited"].shape
        prize, pen = td["cur_total_prize"].clone(), td["cur_total_penalty"].clone()
        visited, i = td["visited"].clone(), td["i"].clone()
        cur = torch.empty(B, dtype=torch.int64, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        done = torch.empty(B, dtype=torch.bool, device=dev)
        status = _lib.deferred_status(dev)
        _lib.call("rl4co_pctsp_step", B, N, _lib.ptr(action.contiguous(), torch.int64),
                  _lib.ptr(td["real_prize"].contiguous(), torch.float32),
                  _lib.ptr(td["penalty"].contiguous(), torch.float32), _lib.ptr(prize), _lib.ptr(pen),
                  _lib.ptr(cur), _lib.ptr(_as_u8(visited)), _lib.ptr(i, torch.int64), _lib.ptr(_as_u8(mask)),
                  _lib.ptr(_as_u8(done)), _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"current_node": cur, "cur_total_prize": prize, "cur_total_penalty": pen, "visited": visited,
                   "i": i, "reward": torch.zeros_like(done), "done": done})
        td.set("action_mask", mask)
        return td

    def _replay_call(self, td, actions, check, want_state):
        locs = td["locs"]
        _need_cuda(locs, "PCTSPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        T = actions.shape[1]
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        st = {}
        if want_state:
            st = {"current_node": torch.empty(B, dtype=torch.int64, device=dev),
                  "cur_total_prize": torch.empty(B, dtype=torch.float32, device=dev),
                  "cur_total_penalty": td["penalty"][..., 1:].sum(-1),      # in/out: the reset state's value
                  "visited": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "action_mask": torch.empty(B, N, dtype=torch.bool, device=dev),
                  "i": torch.empty(B, dtype=torch.int64, device=dev),
                  "done": torch.empty(B, dtype=torch.bool, device=dev)}
        p = lambda k: _lib.ptr(_as_u8(st[k])) if k in st else None  # noqa: E731
        _lib.call("rl4co_pctsp_replay", B, N, T, _lib.ptr(locs.contiguous(), torch.float32),
                  _lib.ptr(td["real_prize"].contiguous(), torch.float32),
                  _lib.ptr(td["penalty"].contiguous(), torch.float32),
                  _lib.ptr(actions.contiguous(), torch.int64), p("current_node"), p("cur_total_prize"),
                  p("cur_total_penalty"), p("visited"), p("action_mask"), p("i"), p("done"),
                  _lib.ptr(reward), int(bool(check)), _lib.ptr(status), _lib.stream_ptr(dev))
        _lib.raise_for_status(status)
        return st, reward

    def replay(self, td, actions, check: Optional[bool] = None):
        """Fused reset-state + T x _step + get_reward.  ``td`` is a reset td."""
        check = self.check_solution if check is None else check
        st, reward = self._replay_call(td, actions, check, True)
        keep = {k: td[k] for k in ("locs", "expected_prize", "real_prize", "penalty", "prize_required")}
        out = TensorDict({**keep, **st, "reward": torch.zeros_like(st["done"])}, batch_size=[actions.shape[0]])
        return out, reward

    def _get_reward(self, td, actions):
        """pctsp/env.py:150-174."""
        if actions.size(-1) == 1:
            assert (actions == 0).all(), "If all length 1 tours, they should be zero"
            return torch.zeros(actions.size(0), dtype=torch.float, device=actions.device)
        return self._replay_call(td, actions, False, False)[1]

    def check_solution_validity(self, td, actions) -> None:
        """pctsp/env.py:176-200: "Duplicates" / "Total prize does not satisfy min total prize"."""
        self._replay_call(td, actions, True, False)


class SPCTSPEnv(PCTSPEnv):
    """rl4co/envs/routing/spctsp/env.py:8-31: the collected prize is the stochastic one, revealed on the visit."""

    name = "spctsp"
    _stochastic = True


# =====================================================================================
class SLAPEnv(RL4COEnvBase):
    """rl4co/envs/warehousing/slap/env.py:17-182."""

    name = "slap"

    def __init__(self, generator: SLAPGenerator = None, generator_params: dict = {},
                 check_solution=False, **kwargs):
        super().__init__(**kwargs)
        if generator is None:
            generator = SLAPGenerator(**{"device": self.device, **generator_params})
        self.generator = generator
        self.check_solution = check_solution

    def _reset(self, td, batch_size=None):
        """slap/env.py:95-129."""
        _need_cuda(td["locs"], "SLAPEnv.reset")
        dev, B = td["locs"].device, int(batch_size[0])
        P, N = td["freq"].shape[-2], td["locs"].shape[1]
        to_choose = torch.arange(P, dtype=torch.float32, device=dev).unsqueeze(0).repeat(B, 1)
        mask = torch.ones(B, N, dtype=torch.bool, device=dev)
        mask[..., 0] = False  # depot is not available (:116)
        return TensorDict({
            "assignment": td["assignment"], "to_choose": to_choose,
            "i": torch.zeros(B, 1, dtype=torch.int64, device=dev),
            "ratio": torch.zeros(td["depot_loc_dist"].shape, device=dev),
            "action_mask": mask, "reward": torch.zeros(B, 1, dtype=torch.float32, device=dev),
        }, batch_size=batch_size)

    def _step(self, td):
        """slap/env.py:38-93 -> rl4co_slap_step (the reference's per-instance python loop
        at :61-62 is one launch here)."""
        action = td["action"]
        _need_cuda(action, "SLAPEnv._step")
        dev, B = action.device, action.shape[0]
        to_choose = td["to_choose"]
        P = td["freq"].shape[-2]
        assignment = td["assignment"].clone()
        mask = td["action_mask"].clone()
        i = td["i"].clone()
        done = torch.empty(B, 1, dtype=torch.bool, device=dev)
        status = _lib.deferred_status(dev)
        if not (to_choose.dim() == 2 and to_choose.stride(1) == 1):
            to_choose = to_choose.contiguous()
        _lib.call("rl4co_slap_step", B, mask.shape[-1], P, _lib.ptr(action.contiguous(), torch.int64),
                  to_choose.data_ptr(), int(to_choose.stride(0)), _lib.ptr(i),
                  _lib.ptr(assignment, torch.int32), _lib.ptr(_as_u8(mask)), _lib.ptr(_as_u8(done)),
                  _lib.ptr(status), _lib.stream_ptr(dev))
        td.update({"assignment": assignment, "to_choose": to_choose[..., 1:], "action_mask": mask,
                   "i": i, "reward": torch.zeros_like(done), "done": done})
        return td

    def get_action_mask(self, td):
        return td["action_mask"]

    def _get_reward(self, td, actions=None):
        """slap/env.py:131-143 -> rl4co_slap_reward (uses td["assignment"], not actions)."""
        locs, assignment, picklist = td["locs"], td["assignment"], td["picklist"]
        _need_cuda(locs, "SLAPEnv.get_reward")
        dev, (B, N, _) = locs.device, locs.shape
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        _lib.call("rl4co_slap_reward", B, N, assignment.shape[-1], picklist.shape[1], picklist.shape[2],
                  _lib.ptr(locs.contiguous(), torch.float32), _lib.ptr(assignment.contiguous(), torch.int32),
                  _lib.ptr(picklist.contiguous(), torch.int64), _lib.ptr(reward), _lib.stream_ptr(dev))
        return reward

    def check_solution_validity(self, td, actions) -> None:
        raise NotImplementedError  # the reference defines none for SLAP (check_solution=False)

    def replay(self, td, actions):
        """Fused reset + P x _step + _get_reward (rl4co_slap_replay)."""
        locs, picklist = td["locs"], td["picklist"]
        _need_cuda(locs, "SLAPEnv.replay")
        dev, (B, N, _) = locs.device, locs.shape
        P = actions.shape[1]
        assignment = torch.empty(B, P, dtype=torch.int32, device=dev)
        mask = torch.empty(B, N, dtype=torch.bool, device=dev)
        i = torch.empty(B, 1, dtype=torch.int64, device=dev)
        done = torch.empty(B, 1, dtype=torch.bool, device=dev)
        reward = torch.empty(B, dtype=torch.float32, device=dev)
        status = _lib.new_status(dev)
        _lib.call("rl4co_slap_replay", B, N, P, picklist.shape[1], picklist.shape[2],
                  _lib.ptr(locs.contiguous(), torch.float32), _lib.ptr(picklist.contiguous(), torch.int64),
                  _lib.ptr(actions.contiguous(), torch.int64), _lib.ptr(assignment), _lib.ptr(_as_u8(mask)),
                  _lib.ptr(i), _lib.ptr(_as_u8(done)), _lib.ptr(reward), _lib.ptr(status),
                  _lib.stream_ptr(dev))
        _lib.raise_for_status(status)
        out = TensorDict({"locs": locs, "picklist": picklist, "assignment": assignment,
                          "action_mask": mask, "i": i, "done": done,
                          "to_choose": torch.empty(B, 0, dtype=torch.float32, device=dev),
                          "reward": torch.zeros_like(done)}, batch_size=[B])
        return out, reward


ENV_REGISTRY = {"tsp": TSPEnv, "atsp": ATSPEnv, "cvrp": CVRPEnv, "sdvrp": SDVRPEnv, "slap": SLAPEnv, "op": OPEnv, "pctsp": PCTSPEnv,
                "spctsp": SPCTSPEnv}


def get_env(env_name: str, *args, **kwargs) -> RL4COEnvBase:
    """rl4co/envs/__init__.py:67-83."""
    env_cls = ENV_REGISTRY.get(env_name)
    if env_cls is None:
        raise ValueError(f"Unknown environment {env_name}. Available environments: {ENV_REGISTRY.keys()}")
    return env_cls(*args, **kwargs)
