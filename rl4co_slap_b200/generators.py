"""Instance generators with the reference's names, kwargs and output keys.

Follows rl4co/envs/common/utils.py:11-92 (Generator, get_sampler),
rl4co/envs/routing/tsp/generator.py:14-60, rl4co/envs/routing/cvrp/generator.py:15-143,
rl4co/envs/warehousing/slap/generator.py:21-155, rl4co/envs/routing/op/generator.py:15-169.

Generators are data preparation, not the hot path: they are vectorised torch code that
runs on whatever ``device`` is asked for (the reference always samples on the CPU; its
SLAP generator is O(B*N) python loops, here it is one broadcast).
"""
from __future__ import annotations

import abc
from typing import Callable, Union

import torch
from torch.distributions import Exponential, Normal, Poisson, Uniform

from .tensordict_compat import TensorDict

# Kool et al. 2019 capacities (cvrp/generator.py:15-30)
CAPACITIES = {10: 20.0, 15: 25.0, 20: 30.0, 30: 33.0, 40: 37.0, 50: 40.0, 60: 43.0, 75: 45.0,
              100: 50.0, 125: 55.0, 150: 60.0, 200: 70.0, 500: 100.0, 1000: 150.0}


class Generator(metaclass=abc.ABCMeta):
    """envs/common/utils.py:11-24: called as ``env.generator(batch_size)``."""

    device = None

    def __init__(self, **kwargs):
        self.kwargs = kwargs

    def __call__(self, batch_size) -> TensorDict:
        batch_size = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        return self._generate(batch_size)

    @abc.abstractmethod
    def _generate(self, batch_size, **kwargs) -> TensorDict:
        raise NotImplementedError


class _DeviceUniform:
    """Uniform(low, high) sampled directly on a device (same distribution as
    torch.distributions.Uniform.sample: low + rand * (high - low))."""

    def __init__(self, low, high, device=None):
        self.low, self.high, self.device = float(low), float(high), device

    def sample(self, shape):
        u = torch.rand(tuple(shape), dtype=torch.float32, device=self.device)
        return self.low + u * (self.high - self.low)


def get_sampler(val_name: str, distribution: Union[int, float, str, type, Callable],
                low: float = 0, high: float = 1.0, device=None, **kwargs):
    """envs/common/utils.py:27-92 (uniform / constant / normal / exponential / poisson /
    center / corner / callable; the cluster / mixture samplers are out of scope)."""
    if isinstance(distribution, (int, float)):
        return _DeviceUniform(distribution, distribution, device)
    if distribution == Uniform or distribution == "uniform":
        return _DeviceUniform(low, high, device)
    if distribution == Normal or distribution in ("normal", "gaussian"):
        assert kwargs.get(val_name + "_mean") is not None, "mean is required for Normal distribution"
        assert kwargs.get(val_name + "_std") is not None, "std is required for Normal distribution"
        return Normal(loc=kwargs[val_name + "_mean"], scale=kwargs[val_name + "_std"])
    if distribution == Exponential or distribution == "exponential":
        assert kwargs.get(val_name + "_rate") is not None, "rate is required for Exponential/Poisson distribution"
        return Exponential(rate=kwargs[val_name + "_rate"])
    if distribution == Poisson or distribution == "poisson":
        assert kwargs.get(val_name + "_rate") is not None, "rate is required for Exponential/Poisson distribution"
        return Poisson(rate=kwargs[val_name + "_rate"])
    if distribution == "center":
        return _DeviceUniform((high - low) / 2, (high - low) / 2, device)
    if distribution == "corner":
        return _DeviceUniform(low, low, device)
    if callable(distribution):
        return distribution(**kwargs)
    raise ValueError(f"Invalid distribution type of {distribution}")


class TSPGenerator(Generator):
    """tsp/generator.py:14-60 -> td{"locs" [B, num_loc, 2] f32}."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0,
                 init_sol_type: str = "random", loc_distribution=Uniform, device=None, **kwargs):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc
        self.init_sol_type = init_sol_type
        self.device = device
        if kwargs.get("loc_sampler") is not None:
            self.loc_sampler = kwargs["loc_sampler"]
        else:
            self.loc_sampler = get_sampler("loc", loc_distribution, min_loc, max_loc, device=device, **kwargs)

    def _generate(self, batch_size) -> TensorDict:
        locs = self.loc_sampler.sample((*batch_size, self.num_loc, 2))
        return TensorDict({"locs": locs}, batch_size=batch_size)


class ATSPGenerator(Generator):
    """atsp/generator.py:14-85 -> td{"cost_matrix" [B,n,n]}: uniform entries, zero diagonal and, with
    ``tmat_class``, the min-plus closure iterated to its fixed point so that the triangle inequality holds
    (MatNet, Kwon et al. 2021).  The closure materialises [b,n,n,n] sums, so it runs over slices of the batch."""

    def __init__(self, num_loc: int = 10, min_dist: float = 0.0, max_dist: float = 1.0, dist_distribution=Uniform,
                 tmat_class: bool = True, device=None, **kwargs):
        self.num_loc, self.min_dist, self.max_dist, self.tmat_class = num_loc, min_dist, max_dist, tmat_class
        self.device = device
        self.dist_sampler = kwargs.get("dist_sampler") or get_sampler("dist", dist_distribution, 0.0, 1.0,
                                                                      device=device, **kwargs)

    def _generate(self, batch_size) -> TensorDict:
        n = self.num_loc
        dms = self.dist_sampler.sample((*batch_size, n, n)) * (self.max_dist - self.min_dist) + self.min_dist
        dms[..., torch.arange(n), torch.arange(n)] = 0
        if self.tmat_class:
            flat = dms.reshape(-1, n, n)
            step = max(1, (1 << 26) // (n * n * n))            # <= 256 MB of fp32 sums at a time
            for c0 in range(0, flat.shape[0], step):
                blk = flat[c0:c0 + step]
                while True:
                    new = (blk[:, :, None, :] + blk[:, None, :, :].transpose(-2, -1)).min(dim=-1)[0]
                    same = bool((new == blk).all())
                    blk = new
                    if same:
                        break
                flat[c0:c0 + step] = blk
            dms = flat.reshape(*batch_size, n, n)
        return TensorDict({"cost_matrix": dms}, batch_size=batch_size)


class CVRPGenerator(Generator):
    """cvrp/generator.py:33-143 -> td{"locs" [B,n,2], "depot" [B,2], "demand" [B,n]
    (already / capacity), "capacity" [B,1]}."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0,
                 loc_distribution=Uniform, depot_distribution=None, min_demand: int = 1,
                 max_demand: int = 10, demand_distribution=Uniform, vehicle_capacity: float = 1.0,
                 capacity: float = None, device=None, **kwargs):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc
        self.min_demand, self.max_demand = min_demand, max_demand
        self.vehicle_capacity = vehicle_capacity
        self.device = device
        self.loc_sampler = kwargs.get("loc_sampler") or get_sampler(
            "loc", loc_distribution, min_loc, max_loc, device=device, **kwargs)
        if kwargs.get("depot_sampler") is not None:
            self.depot_sampler = kwargs["depot_sampler"]
        else:
            self.depot_sampler = (get_sampler("depot", depot_distribution, min_loc, max_loc, device=device, **kwargs)
                                  if depot_distribution is not None else None)
        self.demand_sampler = kwargs.get("demand_sampler") or get_sampler(
            "demand", demand_distribution, min_demand - 1, max_demand - 1, device=device, **kwargs)
        if capacity is None:
            capacity = CAPACITIES.get(num_loc)
        if capacity is None:  # closest table entry (cvrp/generator.py:103-112)
            closest = min(CAPACITIES.keys(), key=lambda x: abs(x - num_loc))
            capacity = CAPACITIES[closest]
        self.capacity = capacity

    def _generate(self, batch_size) -> TensorDict:
        if self.depot_sampler is not None:
            depot = self.depot_sampler.sample((*batch_size, 2))
            locs = self.loc_sampler.sample((*batch_size, self.num_loc, 2))
        else:  # depot = first sampled location (cvrp/generator.py:121-124)
            locs = self.loc_sampler.sample((*batch_size, self.num_loc + 1, 2))
            depot = locs[..., 0, :]
            locs = locs[..., 1:, :]
        demand = self.demand_sampler.sample((*batch_size, self.num_loc))
        demand = (demand.int() + 1).float()
        capacity = torch.full((*batch_size, 1), self.capacity, device=demand.device)
        return TensorDict({"locs": locs, "depot": depot, "demand": demand / self.capacity,
                           "capacity": capacity}, batch_size=batch_size)


# Kool et al. 2019 length budgets (op/generator.py:15-16)
MAX_LENGTHS = {20: 2.0, 50: 3.0, 100: 4.0}


class OPGenerator(Generator):
    """op/generator.py:19-169 -> td{"locs" [B,n,2], "depot" [B,2], "prize" [B,n], "max_length" [B]}.
    ``prize_type``: "dist" (1..100 / 100 by the distance to the depot, the default), "unif", "const"
    (Fischetti et al. 1998, Kool et al. 2019; :133-152)."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0,
                 loc_distribution=Uniform, depot_distribution=None, min_prize: float = 1.0,
                 max_prize: float = 1.0, prize_distribution=Uniform, prize_type: str = "dist",
                 max_length=None, device=None, **kwargs):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc
        self.min_prize, self.max_prize, self.prize_type = min_prize, max_prize, prize_type
        self.device = device
        self.loc_sampler = kwargs.get("loc_sampler") or get_sampler(
            "loc", loc_distribution, min_loc, max_loc, device=device, **kwargs)
        if kwargs.get("depot_sampler") is not None:
            self.depot_sampler = kwargs["depot_sampler"]
        else:
            self.depot_sampler = (get_sampler("depot", depot_distribution, min_loc, max_loc, device=device, **kwargs)
                                  if depot_distribution is not None else None)
        # the reference also builds a prize sampler (:79-88) that _generate never uses; kept only when given
        self.prize_sampler = kwargs.get("prize_sampler")
        if max_length is None:
            max_length = MAX_LENGTHS.get(num_loc)
        if max_length is None:  # closest table entry (:96-103)
            closest = min(MAX_LENGTHS.keys(), key=lambda x: abs(x - num_loc))
            max_length = MAX_LENGTHS[closest]
        self.max_length = max_length

    def _generate(self, batch_size) -> TensorDict:
        if self.depot_sampler is not None:
            depot = self.depot_sampler.sample((*batch_size, 2))
            locs = self.loc_sampler.sample((*batch_size, self.num_loc, 2))
        else:  # depot = first sampled location (:112-116)
            locs = self.loc_sampler.sample((*batch_size, self.num_loc + 1, 2))
            depot = locs[..., 0, :]
            locs = locs[..., 1:, :]
        dev = locs.device
        if self.prize_type == "const":
            prize = torch.ones(*batch_size, self.num_loc, device=dev)
        elif self.prize_type == "unif":
            prize = (1 + torch.randint(0, 100, (*batch_size, self.num_loc), device=dev).float()) / 100
        elif self.prize_type == "dist":
            prize = (depot[..., None, :] - locs).norm(p=2, dim=-1)
            prize = (1 + (prize / prize.max(dim=-1, keepdim=True)[0] * 99).int()).float() / 100
        else:
            raise ValueError(f"Invalid prize_type: {self.prize_type}")
        if not isinstance(self.max_length, torch.Tensor):
            max_length = torch.full((*batch_size,), float(self.max_length), device=dev)
        else:
            max_length = self.max_length.to(dev)
        return TensorDict({"locs": locs, "depot": depot, "prize": prize, "max_length": max_length},
                          batch_size=batch_size)


class PCTSPGenerator(Generator):
    """pctsp/generator.py:19-160 -> td{"locs" [B,n,2], "depot" [B,2], "penalty", "deterministic_prize",
    "stochastic_prize" [B,n]}.  Prizes ~ U(0, 4/n) so that about half of the customers satisfy the constraint
    ``total prize >= 1``; penalties ~ U(0, max_penalty * penalty_factor / n) (Kool et al. 2019)."""

    def __init__(self, num_loc: int = 20, min_loc: float = 0.0, max_loc: float = 1.0, loc_distribution=Uniform,
                 depot_distribution=None, penalty_factor: float = 3.0, prize_required: float = 1.0, device=None,
                 **kwargs):
        self.num_loc, self.min_loc, self.max_loc = num_loc, min_loc, max_loc
        self.penalty_factor, self.prize_required = penalty_factor, prize_required
        self.device = device
        self.loc_sampler = kwargs.get("loc_sampler") or get_sampler(
            "loc", loc_distribution, min_loc, max_loc, device=device, **kwargs)
        if kwargs.get("depot_sampler") is not None:
            self.depot_sampler = kwargs["depot_sampler"]
        else:
            self.depot_sampler = (get_sampler("depot", depot_distribution, min_loc, max_loc, device=device, **kwargs)
                                  if depot_distribution is not None else None)
        self.deterministic_prize_sampler = _DeviceUniform(0.0, 4.0 / num_loc, device)
        self.stochastic_prize_sampler = _DeviceUniform(0.0, 8.0 / num_loc, device)
        max_penalty = kwargs.get("max_penalty")
        if max_penalty is None:
            max_penalty = MAX_LENGTHS.get(num_loc)
        if max_penalty is None:  # closest table entry (:88-97)
            closest = min(MAX_LENGTHS.keys(), key=lambda x: abs(x - num_loc))
            max_penalty = MAX_LENGTHS[closest]
        self.max_penalty = max_penalty * penalty_factor / num_loc
        self.penalty_sampler = _DeviceUniform(0.0, self.max_penalty, device)

    def _generate(self, batch_size) -> TensorDict:
        if self.depot_sampler is not None:
            depot = self.depot_sampler.sample((*batch_size, 2))
            locs = self.loc_sampler.sample((*batch_size, self.num_loc, 2))
        else:
            locs = self.loc_sampler.sample((*batch_size, self.num_loc + 1, 2))
            depot = locs[..., 0, :]
            locs = locs[..., 1:, :]
        penalty = self.penalty_sampler.sample((*batch_size, self.num_loc))
        deterministic_prize = self.deterministic_prize_sampler.sample((*batch_size, self.num_loc))
        stochastic_prize = self.stochastic_prize_sampler.sample((*batch_size, self.num_loc)) * deterministic_prize
        return TensorDict({"locs": locs, "depot": depot, "penalty": penalty,
                           "deterministic_prize": deterministic_prize, "stochastic_prize": stochastic_prize},
                          batch_size=batch_size)


class SLAPGenerator(Generator):
    """slap/generator.py:21-155.  Grid warehouse: location i sits in aisle i // n_locs at
    x = aisle * inter_aisle_dist, y = (i % n_locs) * inter_loc_dist (:67-80); Manhattan
    distance matrix (:51-65); freq ~ U(min_freq, max_freq) [B,P,1] (:138); picklist =
    max_orders x max_products_in_order uniform SKUs (:99-111); assignment = -1 int32."""

    def __init__(self, n_products: int = 20, n_aisles: int = 10, n_locs: int = 10,
                 inter_loc_dist: float = 1, inter_aisle_dist: float = 2.4, min_freq: int = 1,
                 max_freq: int = 20, max_orders: int = 20, max_products_in_order: int = 5,
                 device=None):
        self.n_products, self.n_aisles, self.n_locs = n_products, n_aisles, n_locs
        self.max_orders, self.max_products_in_order = max_orders, max_products_in_order
        self.inter_loc_dist, self.inter_aisle_dist = inter_loc_dist, inter_aisle_dist
        self.min_freq, self.max_freq = min_freq, max_freq
        self.device = device
        self.freq_sampler = _DeviceUniform(min_freq, max_freq, device)

    @staticmethod
    def _get_distance_matrix(locs: torch.Tensor) -> torch.Tensor:
        if locs.dtype not in (torch.float32, torch.float64):
            locs = locs.to(torch.float32)
        return (locs[..., :, None, :] - locs[..., None, :, :]).abs().sum(-1)

    def _calc_coordinates(self, batch_size) -> torch.Tensor:
        total = self.n_aisles * self.n_locs
        idx = torch.arange(total, device=self.device)
        # python-float (double) products rounded once to fp32, like the reference's loop (:76-79)
        x = ((idx // self.n_locs).to(torch.float64) * self.inter_aisle_dist).to(torch.float32)
        y = ((idx % self.n_locs).to(torch.float64) * self.inter_loc_dist).to(torch.float32)
        return torch.stack((x, y), -1).expand(*batch_size, total, 2).contiguous()

    def _create_picklist(self, freq) -> torch.Tensor:
        B = freq.shape[0]
        return torch.randint(0, self.n_products, (B, self.max_orders, self.max_products_in_order),
                             dtype=torch.int64, device=freq.device)

    def _generate(self, batch_size) -> TensorDict:
        freq = self.freq_sampler.sample((*batch_size, self.n_products, 1))
        locs = self._calc_coordinates(batch_size)
        # every instance shares the grid, so the [B,N,N] matrix is a stride-0 view of one [N,N]
        # matrix (the reference materialises it: 40 GB at B = 1 Mi)
        lead = (0,) * len(batch_size)
        dist_mat = self._get_distance_matrix(locs[lead]).expand(*batch_size, -1, -1)
        assignment = torch.full((*batch_size, self.n_products), -1, dtype=torch.int32, device=locs.device)
        picklist = self._create_picklist(freq)
        return TensorDict({"freq": freq, "locs": locs, "dist_mat": dist_mat, "assignment": assignment,
                           "picklist": picklist, "depot_loc_dist": dist_mat[..., 0, :]},
                          batch_size=batch_size)
