"""Side builds of the tensor-core kernel with parts disabled (timing experiments, results are wrong)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rl4co_slap_b200 import build
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for name, flags in (("no_stg", ["-DTC_EXP_NO_STG"]), ("one_mma", ["-DTC_EXP_ONE_MMA"]), ("no_epi", ["-DTC_EXP_NO_EPI"]), ("both", ["-DTC_EXP_ONE_MMA", "-DTC_EXP_NO_EPI"])):
    d = os.path.join(root, "tools", "_exp", name)
    os.makedirs(d, exist_ok=True)
    build.build(force=True, extra_flags=flags, lib=os.path.join(d, "librl4co_exp.so"), obj_dir=os.path.join(d, "obj"))
    print("built", name)
