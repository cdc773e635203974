"""rl4co_slap_b200 - B200-native (sm_100a) environment + Attention-Model decode rollout.

Drop-in for ONE path of j4n1k/rl4co-slap: RL4COEnvBase-style TSP / CVRP / SLAP envs and
AttentionModelPolicy.forward(decode_type=greedy|sampling|multistart_*|evaluate), backed by
hand-written CUDA kernels behind a C ABI (include/rl4co_slap_b200.h).  No CPU fallback.
"""
from . import _lib  # noqa: F401
from .baselines import (REINFORCE_BASELINES_REGISTRY, ExponentialBaseline, MeanBaseline, NoBaseline, RolloutBaseline,  # noqa: F401
                        SharedBaseline, WarmupBaseline, get_reinforce_baseline, global_reward_mean, reinforce_loss,  # noqa: F401
                        reward_stats, shard_bounds)
from .data import (TensorDictDataset, generate_dataset, load_npz_to_tensordict,  # noqa: F401
                   save_tensordict_to_npz)
from .decoding import decode_step, get_log_likelihood, random_policy, rollout  # noqa: F401
from .envs import (ENV_REGISTRY, ATSPEnv, CVRPEnv, OPEnv, PCTSPEnv, RL4COEnvBase, SDVRPEnv, SLAPEnv, SPCTSPEnv, TSPEnv,  # noqa: F401
                   get_env)
from .evaluation import (AugmentationEval, GreedyEval, GreedyMultiStartAugmentEval,  # noqa: F401
                         GreedyMultiStartEval, SamplingEval, evaluate_policy)
from .generators import ATSPGenerator, CVRPGenerator, OPGenerator, PCTSPGenerator, SLAPGenerator, TSPGenerator  # noqa: F401
from .ops import (batchify, gather_by_index, get_num_starts, get_tour_length,  # noqa: F401
                  select_start_nodes, unbatchify)
from .policy import AttentionModelDecoder, AttentionModelPolicy, POMOPolicy  # noqa: F401
from .pomo import POMO, best_of_rollouts  # noqa: F401
from .tensordict_compat import TensorDict  # noqa: F401
from .training import reinforce_step, replay_masks, teacher_forced_log_likelihood  # noqa: F401
from .transforms import StateAugmentation, dihedral_8_augmentation  # noqa: F401

__version__ = "0.1.0"
