"""SURVEY §8(f) rows on the host side: augmentation transforms, npz dataset IO, dataset plumbing and the
rollout-baseline statistics - against golden vectors generated from the unmodified reference
(oracle/gen_golden_eval.py).  CPU, except the baseline statistics (device partial sums, stub policy / env)."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200 import data as D
from rl4co_slap_b200 import transforms as T
from tests.helpers import fixture, t


def test_augmentation_arguments_are_checked_on_the_host():
    """constructor contracts of StateAugmentation (transforms.py:104-133); the arithmetic itself is kernel-only and is
    tested on the GPU (tests/test_gpu_pomo.py)"""
    with pytest.raises(AssertionError, match="num_augment must be 8"):
        T.StateAugmentation(num_augment=4, augment_fn="dihedral8")
    with pytest.raises(ValueError, match="Unknown augment_fn"):
        T.StateAugmentation(augment_fn="nope")
    assert T.get_augment_function("dihedral8") is T.dihedral_8_augmentation_wrapper
    custom = lambda x, n: x      # noqa: E731
    assert T.StateAugmentation(num_augment=3, augment_fn=custom).augmentation is custom
    with pytest.raises(RuntimeError, match="CUDA"):
        T.dihedral_8_augmentation(torch.rand(2, 3, 2))      # no host arithmetic path


def test_npz_roundtrip_and_cvrp_load_data(tmp_path):
    td = R.CVRPGenerator(num_loc=20)(16)
    raw = {"depot": td["depot"], "locs": td["locs"], "demand": td["demand"] * 30.0,
           "capacity": torch.full((16,), 30.0)}
    f = str(tmp_path / "cvrp20.npz")
    D.save_tensordict_to_npz(R.TensorDict(raw, batch_size=[16]), f)
    back = D.load_npz_to_tensordict(f)
    assert back.batch_size == torch.Size([16])
    for k, v in raw.items():
        assert torch.equal(back[k], v), k
    loaded = R.CVRPEnv(device="cpu").load_data(f)        # cvrp/env.py:192-199: demand / capacity
    torch.testing.assert_close(loaded["demand"], td["demand"], rtol=1e-6, atol=1e-7)
    assert D.check_extension("a/b") == "a/b.npz" and D.check_extension("a/b.npz") == "a/b.npz"


def test_generate_npz_follows_the_reference_draw_order():
    d = D.generate_npz("tsp", 7, 10, seed=1234)
    np.random.seed(1234)
    assert np.array_equal(d["locs"], np.random.uniform(size=(7, 10, 2)).astype(np.float32))
    d = D.generate_npz("cvrp", 5, 20, seed=4321)
    np.random.seed(4321)
    assert np.array_equal(d["depot"], np.random.uniform(size=(5, 2)).astype(np.float32))
    assert np.array_equal(d["locs"], np.random.uniform(size=(5, 20, 2)).astype(np.float32))
    assert np.array_equal(d["demand"], np.random.randint(1, 10, size=(5, 20)).astype(np.float32))
    assert (d["capacity"] == 30.0).all()
    assert D.generate_default_datasets_spec() == (("val", 4321), ("test", 1234))


def test_env_dataset_generates_and_loads(tmp_path):
    env = R.TSPEnv(generator_params=dict(num_loc=12), device="cpu")
    ds = env.dataset(batch_size=[33])
    assert len(ds) == 33 and ds[0:5]["locs"].shape == (5, 12, 2)
    assert [b.batch_size[0] for b in ds.batches(16)] == [16, 16, 1]
    f = str(tmp_path / "tsp12.npz")
    D.save_tensordict_to_npz(ds.data, f)
    ds2 = env.dataset(filename=f)
    assert torch.equal(ds2.data["locs"], ds.data["locs"])
    ds3 = env.dataset(batch_size=[4], filename=str(tmp_path / "missing.npz"))    # base.py:262-267: generate instead
    assert len(ds3) == 4


class _StubPolicy(torch.nn.Module):
    """reward = f(locs): enough for the baseline / eval plumbing on the CPU."""

    def __init__(self, rewards):
        super().__init__()
        self.w = torch.nn.Parameter(torch.zeros(1))
        self.rewards = rewards
        self.calls = []

    def forward(self, td, env=None, decode_type=None, **kw):
        idx = td["idx"]
        self.calls.append(int(idx.shape[0]))
        return {"reward": self.rewards.to(idx.device)[idx], "actions": idx[:, None].repeat(1, 3)}


class _StubEnv:
    name = "tsp"

    def reset(self, td):
        return td

    def dataset(self, batch_size):
        n = batch_size[0]
        return D.TensorDictDataset(R.TensorDict({"idx": torch.arange(n)}, batch_size=[n]))


@pytest.mark.gpu
def test_rollout_baseline_ttest_matches_scipy_golden(dev):
    """statistics through rl4co_reward_stats (device partial sums) vs scipy.stats.ttest_rel in the reference run"""
    fx = fixture("rollout_baseline_tsp20")
    bl_r, cand_r = t(fx["baseline"], dev), t(fx["candidate"], dev)
    n = bl_r.shape[0]
    env = _StubEnv()
    bl = R.RolloutBaseline(bl_alpha=0.05)
    bl.setup(_StubPolicy(bl_r), env, batch_size=3, device=dev, dataset_size=n)
    assert bl.policy.calls == [3, 3, 2] and abs(bl.mean - float(bl_r.mean())) < 1e-6
    torch.testing.assert_close(bl.bl_vals, bl_r)
    wrapped = bl.wrap_dataset(env.dataset([n]), env, batch_size=5, device=dev)    # baselines.py:245-258
    torch.testing.assert_close(wrapped.data["extra"], bl_r)
    r, zero = bl.eval(env.dataset([n]).data.to(dev), None, env)
    assert zero == 0 and torch.equal(r, bl_r)
    cand = _StubPolicy(cand_r)
    mean, tstat, p_val, updated = bl.epoch_callback(cand, env, batch_size=4, device=dev, dataset_size=n)
    assert abs(mean - float(cand_r.mean())) < 1e-6
    assert abs(tstat - float(fx["t"])) < 1e-4 * abs(float(fx["t"]))             # scipy.stats.ttest_rel(-cand, -bl)
    assert abs(p_val - float(fx["p_one_sided"])) < 1e-5
    assert updated == (float(fx["p_one_sided"]) < 0.05)
    if updated:
        torch.testing.assert_close(bl.bl_vals, cand_r)
    # a worse candidate never triggers the test (baselines.py:213)
    worse = _StubPolicy(bl.bl_vals - 1.0)
    assert bl.epoch_callback(worse, env, batch_size=4, device=dev, dataset_size=n)[1:] == (None, None, False)
