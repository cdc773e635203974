"""TEST INFRASTRUCTURE ONLY - golden vectors for the SURVEY §8(f) rows from the UNMODIFIED reference.

Runs only in the build container (imports /root/reference through oracle/ref_shim.py):
    python oracle/gen_golden_eval.py
Re-uses the committed policies / instances (tests/golden/policy_{tsp,cvrp}.npz, {tsp,cvrp}20.npz) and stores
* transforms.npz: rl4co/data/transforms.py outputs (dihedral_8_augmentation, symmetric_transform for fixed
  angles, StateAugmentation("dihedral8") of a td);
* eval_{tsp,cvrp}20.npz: (actions, rewards) of the ``_inner`` of GreedyEval, AugmentationEval(dihedral8),
  GreedyMultiStartEval and GreedyMultiStartAugmentEval(dihedral8) of rl4co/tasks/eval.py;
* rollout_baseline_tsp20.npz: RolloutBaseline.rollout rewards, wrap_dataset "extra", and the t / p-value of
  the epoch_callback test (scipy ttest_rel) for a perturbed candidate.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from oracle.gen_golden import OUT, make_env, make_policy, np_  # noqa: E402
from rl4co.data import transforms as T  # noqa: E402
from rl4co.tasks import eval as E  # noqa: E402
from tensordict.tensordict import TensorDict  # noqa: E402


def load_case(env_name, size):
    fx = np.load(os.path.join(OUT, f"{env_name}{size}.npz"))
    pol = make_policy(env_name)
    sd = np.load(os.path.join(OUT, f"policy_{env_name}.npz"))
    pol.load_state_dict({k: torch.from_numpy(sd[k]) for k in sd.files})
    keys = [k[4:] for k in fx.files if k.startswith("gen.") and k != "gen.vehicle_capacity"]
    B = fx["gen.locs"].shape[0]
    td0 = TensorDict({k: torch.from_numpy(fx[f"gen.{k}"]) for k in keys}, batch_size=[B])
    return pol.eval(), make_env(env_name, size), td0


def main():
    torch.manual_seed(7)
    xy = torch.rand(5, 7, 2)
    phi = torch.tensor([0.0, 1.0, 3.5, 7.0, 11.0])[:, None, None]
    data = {"xy": np_(xy), "dihedral8": np_(T.dihedral_8_augmentation(xy)), "phi": np_(phi),
            "symmetric": np_(T.symmetric_transform(xy[..., [0]], xy[..., [1]], phi))}
    td = TensorDict({"locs": xy, "other": torch.arange(5.0)}, batch_size=[5])
    aug = T.StateAugmentation(num_augment=8, augment_fn="dihedral8")(td)
    data["state_aug.locs"], data["state_aug.other"] = np_(aug["locs"]), np_(aug["other"])
    np.savez_compressed(os.path.join(OUT, "transforms.npz"), **data)

    for env_name in ("tsp", "cvrp"):
        pol, env, td0 = load_case(env_name, 20)
        S = 20
        out = {}
        drivers = {"greedy": E.GreedyEval(env, progress=False),
                   "augment": E.AugmentationEval(env, num_augment=8, force_dihedral_8=True, progress=False),
                   "multistart": E.GreedyMultiStartEval(env, num_starts=S, progress=False),
                   "multistart_augment": E.GreedyMultiStartAugmentEval(env, num_starts=S, num_augment=8,
                                                                       force_dihedral_8=True, progress=False)}
        with torch.inference_mode():
            for name, drv in drivers.items():
                drv.env = env
                td = env.reset(td0.clone())
                # the eval drivers call policy(td, ...) without env: give the policy its env by name
                actions, rewards = drv._inner(lambda t, **kw: pol(t, env, **kw), td)
                out[f"{name}.actions"], out[f"{name}.rewards"] = np_(actions), np_(rewards)
                print(env_name, name, float(rewards.mean()))
        out["num_starts"] = np.int64(S)
        np.savez_compressed(os.path.join(OUT, f"eval_{env_name}20.npz"), **out)

    # RolloutBaseline: rollout / t-test numbers
    from scipy.stats import ttest_rel
    pol, env, td0 = load_case("tsp", 20)
    with torch.inference_mode():
        bl = pol(env.reset(td0.clone()), env, decode_type="greedy")["reward"]
        cand_pol = make_policy("tsp")
        torch.manual_seed(99)
        sd = {k: v + 0.02 * torch.randn_like(v) if v.dtype.is_floating_point else v for k, v in pol.state_dict().items()}
        cand_pol.load_state_dict(sd)
        cand = cand_pol.eval()(env.reset(td0.clone()), env, decode_type="greedy")["reward"]
    t, p = ttest_rel(-np_(cand), -np_(bl))
    np.savez_compressed(os.path.join(OUT, "rollout_baseline_tsp20.npz"), baseline=np_(bl), candidate=np_(cand),
                        t=np.float64(t), p_one_sided=np.float64(p / 2),
                        **{f"cand.{k}": np_(v) for k, v in sd.items()})
    print("t", t, "p/2", p / 2, "means", float(cand.mean()), float(bl.mean()))


if __name__ == "__main__":
    main()
