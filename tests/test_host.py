"""CPU: the C-ABI library loads and exports every declared symbol; host-side logic
(generators, batchify, TensorDict compat, policy weight folding, sharding + reductions over
a 2-rank gloo group); the product refuses CPU tensors instead of falling back."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib
from tests.helpers import fixture, gen_td, load_policy, oracle_weights, t

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "rl4co_slap_b200.h")).read()
    declared = set(re.findall(r"\b(rl4co_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTED_SYMBOLS), declared ^ set(_lib.EXPORTED_SYMBOLS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.rl4co_abi_version() == 4


def test_args_struct_matches_header():
    header = open(os.path.join(ROOT, "include", "rl4co_slap_b200.h")).read()
    body = header.split("typedef struct rl4co_am_rollout_args {")[1].split("} rl4co_am_rollout_args;")[0]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            names += [n.strip().lstrip("*") for n in re.sub(r"^(const\s+)?\w+\s+", "", decl).split(",")]
    assert names == [f[0] for f in _lib.AmRolloutArgs._fields_]


def test_host_argument_error_is_reported_not_fatal():
    lib = _lib.load()
    rc = lib.rl4co_tsp_reset(-1, 0, None, None, None, None, None, None, None, None)
    assert rc != 0 and b"bad sizes" in lib.rl4co_last_error()


def test_no_cpu_fallback():
    env = R.TSPEnv(generator_params=dict(num_loc=5), device="cpu")
    td = env.generator(2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        env.reset(td)
    pol = R.AttentionModelPolicy(env_name="tsp", num_encoder_layers=1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pol(td, env, decode_type="greedy")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        R.decode_step(torch.zeros(2, 5), torch.ones(2, 5, dtype=torch.bool))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        R.reward_stats(torch.zeros(4))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "rl4co_slap_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), fn


def test_generators_shapes_and_ranges():
    td = R.TSPGenerator(num_loc=50)(7)
    assert td["locs"].shape == (7, 50, 2) and td["locs"].dtype == torch.float32
    assert 0 <= td["locs"].min() and td["locs"].max() < 1
    g = R.CVRPGenerator(num_loc=100)
    td = g([5])
    assert td["locs"].shape == (5, 100, 2) and td["depot"].shape == (5, 2)
    assert td["demand"].shape == (5, 100) and td["capacity"].shape == (5, 1)
    assert g.capacity == 50.0
    d = td["demand"] * 50.0
    assert torch.allclose(d, d.round()) and d.min() >= 1 and d.max() <= 9
    assert R.CVRPGenerator(num_loc=33).capacity == 33.0
    td = R.SLAPGenerator(n_aisles=5, n_locs=10)(3)
    fx = fixture("slap50")
    assert np.array_equal(td["locs"].numpy(), fx["gen.locs"][:3])  # same grid as the reference
    assert np.array_equal(td["depot_loc_dist"].numpy(), fx["gen.depot_loc_dist"][:3])
    assert td["picklist"].shape == (3, 20, 5) and td["picklist"].dtype == torch.int64
    assert td["assignment"].dtype == torch.int32 and (td["assignment"] == -1).all()
    assert td["freq"].shape == (3, 20, 1) and td["dist_mat"].shape == (3, 50, 50)


@pytest.mark.parametrize("shape", [(2,), (2, 2), (2, 2, 2)])
def test_batchify_roundtrip(shape):
    """the reference's own property test (tests/test_utils.py:10-29)"""
    x = torch.rand(4, 5)
    y = R.unbatchify(R.batchify(x, shape), shape)
    assert torch.equal(y[(slice(None),) + (0,) * len(shape)], x)
    td = R.TensorDict({"a": torch.rand(4, 3), "b": torch.rand(4)}, batch_size=[4])
    back = R.unbatchify(R.batchify(td, shape), shape)
    assert torch.equal(back["a"][(slice(None),) + (0,) * len(shape)], td["a"])


def test_batchify_is_start_major():
    x = torch.arange(3)
    assert R.batchify(x, 2).tolist() == [0, 1, 2, 0, 1, 2]
    env = R.CVRPEnv(generator_params=dict(num_loc=4), device="cpu")
    td = R.TensorDict({"action_mask": torch.ones(3, 5, dtype=torch.bool)}, batch_size=[3])
    assert env.get_num_starts(td) == 4
    assert env.select_start_nodes(td, 4).tolist() == [1, 1, 1, 2, 2, 2, 3, 3, 3, 4, 4, 4]
    tsp = R.TSPEnv(generator_params=dict(num_loc=5), device="cpu")
    assert tsp.select_start_nodes(td, 2).tolist() == [0, 0, 0, 1, 1, 1]


def test_state_dict_names_and_folded_weights():
    """reference checkpoints load unchanged; folded tables reproduce the reference algebra"""
    for name in ("tsp", "cvrp", "slap"):
        pol = load_policy(name, "cpu")
        W, q0, wcap, _ = pol.decoder.folded_weights()
        ow = oracle_weights(name)
        E = 128
        h = torch.randn(6, E)
        o = torch.randn(E)
        tab = h @ W.t()
        lk = h @ ow["Wnode"][2 * E:].t()
        ref_logit = lk @ (ow["Wout"] @ o)
        torch.testing.assert_close(tab[:, 2 * E:3 * E] @ o, ref_logit, rtol=1e-4, atol=1e-4)
        if name == "tsp":
            ctx = torch.cat([h[0], h[3]]) @ ow["Wctx"].t()
            torch.testing.assert_close(tab[0, 4 * E:5 * E] + tab[3, 3 * E:4 * E], ctx, rtol=1e-4, atol=1e-5)
            torch.testing.assert_close(q0, ow["Wctx"] @ ow["Wph"], rtol=1e-5, atol=1e-5)
        if name == "cvrp":
            ctx = torch.cat([h[2], torch.tensor([0.25])]) @ ow["Wctx"].t()
            torch.testing.assert_close(tab[2, 3 * E:4 * E] + 0.25 * wcap, ctx, rtol=1e-4, atol=1e-5)


def test_encoder_matches_reference_hidden():
    """PyTorch encoder (stays PyTorch) reproduces the reference's node embeddings"""
    for case, name in (("tsp20", "tsp"), ("cvrp20", "cvrp"), ("slap50", "slap")):
        fx = fixture(case)
        pol = load_policy(name, "cpu")
        td = gen_td(name, fx)
        if name == "cvrp":
            td = R.TensorDict({"locs": torch.cat((td["depot"][:, None], td["locs"]), 1), "demand": td["demand"]},
                              batch_size=td.batch_size)
        with torch.inference_mode():
            h, _ = pol.encoder(td)
        np.testing.assert_allclose(h.numpy(), fx["hidden"], rtol=1e-4, atol=1e-5)


def test_shard_bounds_cover_everything():
    for total, world in ((10, 3), (1 << 20, 8), (5, 8)):
        spans = [R.shard_bounds(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from rl4co_slap_b200 import baselines as B
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
torch.manual_seed(7)
reward = -torch.rand(1001, dtype=torch.float32) * 10          # the global reward vector (same on both ranks)
ll = -torch.rand(1001, dtype=torch.float32)
lo, hi = B.shard_bounds(1001, rank, 2)
r = reward[lo:hi].double()
stats = torch.stack((r.sum(), (r * r).sum(), torch.tensor(float(hi - lo), dtype=torch.float64)))
mean, var, n = B.finalize_stats(stats).tolist()
assert abs(mean - reward.double().mean().item()) < 1e-12, (mean,)
assert abs(var - reward.double().var().item()) < 1e-9 and n == 1001
adv = (reward - reward.mean())[lo:hi].double()
part = torch.stack(((adv * ll[lo:hi].double()).sum(), torch.tensor(float(hi - lo), dtype=torch.float64)))
loss = B.reinforce_loss(part).item()
want = -((reward - reward.mean()).double() * ll.double()).mean().item()
assert abs(loss - want) < 1e-12, (loss, want)
d = torch.randn(1001, dtype=torch.float64)[lo:hi] + 0.1
tstat, n = B.paired_ttest_finalize(torch.stack((d.sum(), (d * d).sum(), torch.tensor(float(hi - lo), dtype=torch.float64))))
assert n == 1001 and tstat == tstat

# RolloutBaseline over a sharded dataset: every rank rolls out its own slice, the all-reduced t-test equals
# scipy.stats.ttest_rel on the FULL data (baselines.py:202-224)
import scipy.stats
def cpu_stats(reward, stats=None):           # host stand-in of rl4co_reward_stats (the kernel needs a GPU)
    r = reward.double().reshape(-1)
    s = torch.stack((r.sum(), (r * r).sum(), torch.tensor(float(r.numel()), dtype=torch.float64)))
    return s if stats is None else stats.add_(s)
B.reward_stats = cpu_stats
from rl4co_slap_b200 import TensorDict
data = TensorDict({"locs": torch.rand(1001, 5, 2)}, batch_size=[1001])      # same seed on both ranks
class FakeEnv:
    def reset(self, td): return td
    def dataset(self, batch_size): return data
class FakePolicy(torch.nn.Module):
    def __init__(self, shift):
        super().__init__(); self.shift = shift
    def forward(self, td, env, decode_type="greedy"):
        return {"reward": -td["locs"].sum((1, 2)) + self.shift * td["locs"][:, 0, 0]}
rb = B.RolloutBaseline(bl_alpha=0.05)
rb._update_policy(FakePolicy(0.0), FakeEnv(), batch_size=100, device="cpu", dataset=data)
assert rb.bl_vals.numel() == hi - lo, rb.bl_vals.shape
full_bl, full_c = FakePolicy(0.0)(data, None)["reward"], FakePolicy(0.5)(data, None)["reward"]
assert abs(rb.mean - full_bl.double().mean().item()) < 1e-9
cand_mean, tstat, p_val, updated = rb.epoch_callback(FakePolicy(0.5), FakeEnv(), batch_size=100, device="cpu", epoch=0)
want = scipy.stats.ttest_rel(-full_c.double().numpy(), -full_bl.double().numpy())
assert abs(cand_mean - full_c.double().mean().item()) < 1e-9
assert abs(tstat - want.statistic) < 1e-6 * abs(want.statistic), (tstat, want.statistic)
assert abs(p_val - want.pvalue / 2) <= 1e-6 * max(want.pvalue, 1e-300) + 1e-300, (p_val, want.pvalue)
assert updated and abs(rb.mean - full_c.double().mean().item()) < 1e-9
wrapped = rb.wrap_dataset(data, FakeEnv(), batch_size=100, device="cpu")
assert wrapped.data["extra"].numel() == 1001          # wrap_dataset covers every local row, not a shard
dist.barrier(); dist.destroy_process_group()
print("ok", rank)
"""


def test_sharded_reductions_world2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


def test_gumbel_uniform_is_strictly_inside_unit_interval():
    """csrc/rollout_util.cuh gumbel4: u = ((w >> 9) + 0.5) * 2^-23 in fp32 is exact and in (0,1) for every 32-bit
    Philox word, so the Gumbel noise -ln(-ln u) is finite (the 24-bit form rounded w = 2^32 - 1 to u = 1 -> +inf)"""
    w = np.array([0, 1, 511, 512, 2 ** 31, 2 ** 32 - 512, 2 ** 32 - 1], dtype=np.uint64)
    k = (w >> np.uint64(9)).astype(np.float32)
    u = (k + np.float32(0.5)) * np.float32(1.0 / 8388608.0)
    assert u.dtype == np.float32 and (u > 0).all() and (u < 1).all()
    assert np.array_equal(u.astype(np.float64), ((w >> np.uint64(9)).astype(np.float64) + 0.5) / 8388608.0)   # exact
    g = -np.log(-np.log(u.astype(np.float64)))
    assert np.isfinite(g).all()
    old = (np.float32(2 ** 24 - 1) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)
    assert old == np.float32(1.0)       # what the fix removes
