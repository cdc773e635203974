"""TEST INFRASTRUCTURE ONLY - golden vectors of the POMO val / test epilogue and of multisample
decoding, from the UNMODIFIED reference (imports /root/reference through oracle/ref_shim.py; build container only):

    python oracle/gen_golden_pomo.py

Writes tests/golden/pomo_step_{tsp,cvrp}20.npz:
* ``aug8.*``   POMO(env, policy, num_augment=8, augment_fn="dihedral8", num_starts=S).shared_step(batch, 0, "test")
               (rl4co/models/zoo/pomo/model.py:87-144): reward, actions [B,8,S,T], max_reward [B,8],
               best_multistart_actions [B,8,T], max_aug_reward [B], best_aug_actions [B,T];
* ``noaug.*``  the same with num_augment=0: reward, max_reward [B], best_multistart_actions [B,T];
* ``multisample_greedy.*``  policy(td, env, decode_type="greedy", num_starts=3, multisample=True, return_actions=True)
               (rl4co/utils/decoding.py:271-308): the td is tiled 3x and NO first action is forced, so every copy is the
               plain greedy tour - the deterministic witness of what ``multisample`` means in this fork.
The Lightning side of the module (log_metrics) is replaced by a capture of the ``out`` dictionary it is handed.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from oracle.gen_golden import OUT, np_  # noqa: E402
from oracle.gen_golden_eval import load_case  # noqa: E402
from rl4co.models.zoo.pomo.model import POMO  # noqa: E402


def shared_step_out(env, pol, td0, **kw):
    model = POMO(env, policy=pol, **kw)
    seen = {}
    model.log_metrics = lambda out, phase, dataloader_idx=None: (seen.update(out), {})[1]
    with torch.inference_mode():
        model.shared_step(td0.clone(), 0, "test")
    return seen


def main():
    S = 20
    for env_name in ("tsp", "cvrp"):
        pol, env, td0 = load_case(env_name, 20)
        data = {"num_starts": np.int64(S)}
        for tag, kw in (("aug8", dict(num_augment=8, augment_fn="dihedral8", num_starts=S)),
                        ("noaug", dict(num_augment=0, num_starts=S))):
            out = shared_step_out(env, pol, td0, **kw)
            for k, v in out.items():
                if isinstance(v, torch.Tensor):
                    data[f"{tag}.{k}"] = np_(v)
            print(env_name, tag, {k: tuple(v.shape) for k, v in out.items() if isinstance(v, torch.Tensor)})
        with torch.inference_mode():
            ms = pol(env.reset(td0.clone()), env, decode_type="greedy", num_starts=3, multisample=True, return_actions=True)
            plain = pol(env.reset(td0.clone()), env, decode_type="greedy", return_actions=True)
        B = plain["actions"].shape[0]
        assert ms["actions"].shape[0] == 3 * B
        assert all(torch.equal(ms["actions"][i * B:(i + 1) * B], plain["actions"]) for i in range(3)), \
            "multisample forced a first action after all"
        data["multisample_greedy.actions"], data["multisample_greedy.reward"] = np_(ms["actions"]), np_(ms["reward"])
        data["multisample_greedy.ll"] = np_(ms["log_likelihood"])
        np.savez_compressed(os.path.join(OUT, f"pomo_step_{env_name}20.npz"), **data)


if __name__ == "__main__":
    main()
