"""SURVEY 8f rank 3 on the GPU: the augmentation kernels on the way into the encoder batch, the best-of-starts /
best-of-augmentations kernel with its tour gathers, ``POMO.shared_step`` with the reference's output keys, and the
``multisample`` decode option - against golden vectors of the UNMODIFIED reference (oracle/gen_golden_eval.py,
oracle/gen_golden_pomo.py)."""
import numpy as np
import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200 import transforms as T
from rl4co_slap_b200.pomo import POMO, best_of_rollouts
from tests.helpers import fixture, gen_td, load_policy, t

pytestmark = pytest.mark.gpu


def test_dihedral8_and_symmetric_kernels_match_reference(dev):
    fx = fixture("transforms")
    xy = t(fx["xy"], dev)
    assert np.array_equal(T.dihedral_8_augmentation(xy).cpu().numpy(), fx["dihedral8"])       # bit exact
    out = T.symmetric_transform(xy, t(fx["phi"], dev).reshape(-1))
    np.testing.assert_allclose(out.cpu().numpy(), fx["symmetric"], rtol=0, atol=3e-7)       # cosf / sinf: <= 2 ulp
    with pytest.raises(RuntimeError, match="CUDA"):
        T.dihedral_8_augmentation(xy.cpu())


def test_state_augmentation_dihedral8_matches_reference(dev):
    fx = fixture("transforms")
    td = R.TensorDict({"locs": t(fx["xy"], dev), "other": torch.arange(5.0, device=dev)}, batch_size=[5])
    aug = T.StateAugmentation(num_augment=8, augment_fn="dihedral8")(td)
    assert aug.batch_size == torch.Size([40])
    assert np.array_equal(aug["locs"].cpu().numpy(), fx["state_aug.locs"])
    assert np.array_equal(aug["other"].cpu().numpy(), fx["state_aug.other"])
    assert td["locs"].shape == (5, 7, 2)                                                    # the input td is untouched


def test_symmetric_augmentation_keeps_first_copy_and_distances(dev):
    torch.manual_seed(0)
    xy = torch.rand(6, 9, 2, device=dev)
    rep = R.batchify(xy, 4)
    aug = T.symmetric_augmentation(rep, 4)
    torch.testing.assert_close(aug[:6], xy, rtol=0, atol=6e-8)      # phi = 0 for the first copy (transforms.py:86-87)
    d0 = (rep[:, :, None] - rep[:, None]).norm(dim=-1)
    d1 = (aug[:, :, None] - aug[:, None]).norm(dim=-1)
    torch.testing.assert_close(d0, d1, rtol=1e-5, atol=1e-6)        # rotations / reflections are isometries
    assert not torch.allclose(aug[6:], rep[6:])
    td = R.TensorDict({"locs": xy}, batch_size=[6])
    out = T.StateAugmentation(num_augment=4, augment_fn="symmetric")(td)
    assert out["locs"].shape == (24, 9, 2)
    torch.testing.assert_close(out["locs"][:6], xy, rtol=0, atol=6e-8)


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_pomo_best_kernel_on_the_reference_rollouts(name, dev):
    """the epilogue kernel alone: fed the reference's own rewards and tours (policy output order), it must return the
    reference's max_reward / best_multistart_actions / max_aug_reward / best_aug_actions bit for bit"""
    fx = fixture(f"pomo_step_{name}20")
    S = int(fx["num_starts"])
    acts = t(fx["aug8.actions"], dev)                              # [B, A, S, T] = unbatchify(actions, (A, S))
    B, A, _, Tn = acts.shape
    flat = acts.permute(2, 1, 0, 3).reshape(-1, Tn).contiguous()    # back to rollout (s*A + a)*B + b
    best = best_of_rollouts(t(fx["aug8.reward"], dev), flat, B, A, S)
    assert np.array_equal(best["max_reward"].cpu().numpy(), fx["aug8.max_reward"])
    assert np.array_equal(best["best_multistart_actions"].cpu().numpy(), fx["aug8.best_multistart_actions"])
    assert np.array_equal(best["max_aug_reward"].cpu().numpy(), fx["aug8.max_aug_reward"])
    assert np.array_equal(best["best_aug_actions"].cpu().numpy(), fx["aug8.best_aug_actions"])
    r = t(fx["aug8.reward"], dev).view(S, A, B)
    assert torch.equal(best["best_start"], r.permute(2, 1, 0).max(-1).indices)
    # no augmentation: [B, S]
    acts = t(fx["noaug.actions"], dev)
    flat = acts.permute(1, 0, 2).reshape(-1, acts.shape[-1]).contiguous()
    best = best_of_rollouts(t(fx["noaug.reward"], dev), flat, B, 0, S)
    assert np.array_equal(best["max_reward"][:, 0].cpu().numpy(), fx["noaug.max_reward"])
    assert np.array_equal(best["best_multistart_actions"][:, 0].cpu().numpy(), fx["noaug.best_multistart_actions"])
    # ties go to the lowest index (torch.max), ragged block sizes
    rew = torch.zeros(3 * 2 * 301, device=dev)
    tie = best_of_rollouts(rew, None, 301, 2, 3)
    assert (tie["best_start"] == 0).all() and (tie["best_aug"] == 0).all()


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_pomo_shared_step_matches_reference(name, dev):
    """POMO.shared_step (test phase, dihedral-8, S = 20) end to end on the GPU - encoder included, so a greedy near
    tie may flip a tour: the keys and shapes are the reference's, rewards agree on (almost) every entry, and the
    returned best tours are exactly the ones the returned rewards belong to"""
    fx, gold = fixture(f"{name}20"), fixture(f"pomo_step_{name}20")
    S = int(gold["num_starts"])
    env, pol = R.get_env(name, device=dev), load_policy(name, dev)
    model = POMO(env, policy=pol, num_augment=8, augment_fn="dihedral8", num_starts=S)
    assert pol.test_decode_type == "multistart_greedy"
    out = model.shared_step(gen_td(name, fx, dev), 0, "test")
    B = fx["gen.locs"].shape[0]
    assert set(out) >= {"reward", "log_likelihood", "actions", "max_reward", "best_multistart_actions",
                        "max_aug_reward", "best_aug_actions"}
    assert out["actions"].shape[:3] == (B, 8, S) and out["max_reward"].shape == (B, 8)
    assert out["best_multistart_actions"].shape[:2] == (B, 8) and out["max_aug_reward"].shape == (B,)
    for key, frac in (("reward", 0.9), ("max_reward", 0.9), ("max_aug_reward", 0.85)):
        got, want = out[key].cpu().numpy(), gold[f"aug8.{key}"]
        close = np.abs(got - want) <= 1e-4 * np.abs(want)
        assert close.mean() >= frac, (key, close.mean())
    # internal consistency is exact: the kernel's picks against torch on the SAME rollouts
    r = out["reward"].view(S, 8, B).permute(2, 1, 0)
    assert torch.equal(out["max_reward"], r.max(-1).values)
    assert torch.equal(out["max_aug_reward"], r.max(-1).values.max(-1).values)
    a_idx = r.max(-1).values.max(-1).indices
    s_idx = r.max(-1).indices.gather(1, a_idx[:, None])[:, 0]
    assert torch.equal(out["best_aug_actions"], out["actions"][torch.arange(B, device=dev), a_idx, s_idx])
    assert torch.equal(out["best_multistart_actions"],
                       out["actions"].gather(2, r.max(-1).indices[:, :, None, None].expand(-1, -1, 1, out["actions"].shape[-1]))[:, :, 0])
    # no augmentation (validation with num_augment = 0): [B, S] keys
    model0 = POMO(env, policy=pol, num_augment=0, num_starts=S)
    out0 = model0.shared_step(gen_td(name, fx, dev), 0, "val")
    assert out0["max_reward"].shape == (B,) and out0["best_multistart_actions"].dim() == 2 and "max_aug_reward" not in out0
    close = np.abs(out0["max_reward"].cpu().numpy() - gold["noaug.max_reward"]) <= 1e-4 * np.abs(gold["noaug.max_reward"])
    assert close.mean() >= 0.85
    # train phase: loss of the reference's formula with the shared baseline, max_reward [B]
    torch.manual_seed(1)
    tr = model0.shared_step(gen_td(name, fx, dev), 0, "train", optimizer=None)
    r2 = R.unbatchify(tr["reward"], S)
    adv = (r2 - r2.mean(1, keepdim=True)).t().reshape(-1)
    torch.testing.assert_close(tr["loss"], -(adv * tr["log_likelihood"]).mean(), rtol=1e-5, atol=1e-6)
    assert torch.equal(tr["max_reward"], r2.max(1).values)


@pytest.mark.parametrize("name", ("tsp", "cvrp"))
def test_multisample_decoding(name, dev):
    """multisample (utils/decoding.py:271-308): the batch is tiled and NO first action is forced.  Greedy: every copy
    is the plain greedy tour (the reference fixture says the same); sampling: copies differ, first actions are drawn
    from the policy (not forced to s % num_loc), ll = evaluate-mode ll, select_best returns the best copy"""
    fx, gold = fixture(f"{name}20"), fixture(f"pomo_step_{name}20")
    env, pol = R.get_env(name, device=dev), load_policy(name, dev)
    td = env.reset(gen_td(name, fx, dev))
    h = t(fx["hidden"], dev)
    B = h.shape[0]
    out = pol(td, env, decode_type="greedy", num_starts=3, multisample=True, return_actions=True, hidden=h)
    assert np.array_equal(out["actions"].cpu().numpy(), gold["multisample_greedy.actions"])
    np.testing.assert_allclose(out["reward"].cpu().numpy(), gold["multisample_greedy.reward"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out["log_likelihood"].cpu().numpy(), gold["multisample_greedy.ll"], rtol=1e-5, atol=1e-5)
    torch.manual_seed(4)
    S = 64
    smp = pol(td, env, decode_type="sampling", num_starts=S, multisample=True, return_actions=True, hidden=h,
              select_start_nodes_fn=lambda *a: None)
    assert smp["actions"].shape[0] == S * B and smp["reward"].shape == (S * B,)
    first = smp["actions"][:, 0].view(S, B)
    forced = (torch.arange(S, device=dev) % (h.shape[1] - (name != "tsp")) + (name != "tsp"))[:, None].expand(S, B)
    assert not torch.equal(first, forced), "the first action was forced like multistart"
    assert smp["actions"].view(S, B, -1)[:, 0].unique(dim=0).shape[0] > S // 4
    ev = pol(R.batchify(td, S), env, actions=smp["actions"], hidden=R.batchify(h, S))
    torch.testing.assert_close(ev["log_likelihood"], smp["log_likelihood"], rtol=1e-5, atol=2e-5)
    torch.manual_seed(4)
    best = pol(td, env, decode_type="sampling", num_starts=S, multisample=True, return_actions=True, hidden=h,
               select_best=True)
    assert best["reward"].shape == (B,)
    torch.testing.assert_close(best["reward"], R.unbatchify(smp["reward"], S).max(1).values)
