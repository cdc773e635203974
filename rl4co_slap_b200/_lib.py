"""ctypes binding of include/rl4co_slap_b200.h.

There is NO CPU fallback: if the shared library is missing, or a tensor is not on a CUDA
device, calls raise.  `load()` only dlopens the library (works without a GPU, used by the
CPU test-suite to check the exported symbols).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librl4co_slap_b200.so")

ST_INVALID_TOUR, ST_OVER_CAPACITY, ST_INFEASIBLE_ACTION = 1, 2, 4
ST_NAN_LOGITS, ST_NEG_INF_LOGP, ST_ACTION_RANGE = 8, 16, 32
ST_DUPLICATES, ST_OVER_LENGTH, ST_PRIZE_SHORT = 64, 128, 256
ST_DEPOT_TWICE, ST_DEMAND_LEFT = 512, 1024
# messages are the reference's assertion texts (tsp/env.py:173, cvrp/env.py:175,190,
# utils/decoding.py:59,379, models/nn/attention.py:285)
STATUS_MESSAGES = [
    (ST_INVALID_TOUR, "Invalid tour"),
    (ST_OVER_CAPACITY, "Used more than capacity"),
    (ST_INFEASIBLE_ACTION, "infeasible action selected"),
    (ST_NAN_LOGITS, "Logits contain NaNs"),
    (ST_NEG_INF_LOGP, "Logprobs should not be -inf, check sampling procedure!"),
    (ST_ACTION_RANGE, "index out of range in action"),
    (ST_DUPLICATES, "Duplicates"),                       # op/env.py:199
    (ST_OVER_LENGTH, "Max length exceeded"),             # op/env.py:213
    (ST_PRIZE_SHORT, "Total prize does not satisfy min total prize"),   # pctsp/env.py:200
    (ST_DEPOT_TWICE, "Cannot visit depot twice if any nonzero demand"),   # sdvrp/env.py:157
    (ST_DEMAND_LEFT, "All demand must be satisfied"),                     # sdvrp/env.py:164
]
ENV_IDS = {"tsp": 0, "cvrp": 1, "slap": 2, "op": 3, "pctsp": 4, "spctsp": 4, "sdvrp": 5}
MODE_GREEDY, MODE_SAMPLING, MODE_EVALUATE = 0, 1, 2
STATIC_TABLE_BYTES = 32 << 20       # RL4CO_STATIC_TABLE_BYTES: scratch of rl4co_am_rollout_static

_p = C.c_void_p
_i64, _i32, _f32, _u64 = C.c_int64, C.c_int32, C.c_float, C.c_uint64


class AmRolloutArgs(C.Structure):
    _fields_ = [
        ("env", _i32), ("mode", _i32), ("B", _i64), ("S", _i32), ("multistart", _i32),
        ("num_loc", _i32), ("N", _i32), ("T", _i32), ("P", _i32), ("O", _i32), ("K", _i32),
        ("n_tables", _i32), ("temperature", _f32), ("tanh_clipping", _f32), ("mask_logits", _i32),
        ("tables", _p), ("gctx", _p), ("q0", _p), ("wcap", _p), ("locs", _p), ("demand", _p),
        ("vehicle_capacity", _p), ("picklist", _p), ("actions_in", _p),
        ("seed", _u64), ("offset", _u64),
        ("actions", _p), ("logp", _p), ("logp_all", _p), ("ll", _p), ("reward", _p),
        ("steps", _p), ("assignment", _p), ("status", _p),
        ("prize", _p), ("max_length", _p), ("penalty", _p), ("prize_required", _p), ("wdyn", _p),
        ("start_nodes", _p), ("h", _p), ("w_first", _p),
    ]


_SIGNATURES = {
    "rl4co_abi_version": ([], C.c_int),
    "rl4co_last_error": ([], C.c_char_p),
    "rl4co_tsp_reset": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_tsp_step": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_tsp_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_atsp_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_cvrp_action_mask": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_cvrp_step": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_cvrp_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_op_action_mask": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_op_step": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_op_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_pctsp_action_mask": ([_i64, _i32, _p, _p, _p, _p], C.c_int),
    "rl4co_pctsp_step": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_pctsp_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_sdvrp_action_mask": ([_i64, _i32, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_sdvrp_step": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_sdvrp_replay": ([_i64, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_slap_step": ([_i64, _i32, _i32, _p, _p, _i64, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_slap_reward": ([_i64, _i32, _i32, _i32, _i32, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_slap_replay": ([_i64, _i32, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_am_project": ([_i64, _i32, _p, _p, _p, _p], C.c_int),
    "rl4co_am_weight_image_floats": ([_i32], C.c_int64),
    "rl4co_am_pack_weights": ([_i32, _p, _p, _p], C.c_int),
    "rl4co_am_project_tc": ([_i64, _i32, _p, _p, _p, _p], C.c_int),
    "rl4co_linear_pack": ([_i32, _i32, _p, _p, _p], C.c_int),
    "rl4co_linear_narrow": ([_i64, _i32, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_linear_tc": ([_i64, _i32, _i32, _p, _p, _p, _i32, _p, _p], C.c_int),
    "rl4co_linear_tc_ex": ([_i64, _i32, _i32, _p, _p, _p, _i32, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_encoder_attention": ([_i64, _i32, _p, _p, _p], C.c_int),
    "rl4co_encoder_attention_tc": ([_i64, _i32, _p, _p, _p], C.c_int),
    "rl4co_set_attention_impl": ([_i32], C.c_int),
    "rl4co_am_graph_context": ([_i64, _i32, _p, _p, _p, _p], C.c_int),
    "rl4co_am_rollout": ([C.POINTER(AmRolloutArgs), _p], C.c_int),
    "rl4co_am_rollout_static": ([C.POINTER(AmRolloutArgs), _p, _p, _p], C.c_int),
    "rl4co_am_evaluate_backward_max_nodes": ([], C.c_int),
    "rl4co_am_evaluate_backward": ([C.POINTER(AmRolloutArgs), _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_set_rollout_impl": ([C.c_int], C.c_int),
    "rl4co_decode_step": ([_i64, _i32, _p, _p, _i32, _f32, _f32, _i32, _p, _u64, _u64, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_reward_stats": ([_i64, _p, _p, _p], C.c_int),
    "rl4co_shared_baseline": ([_i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
    "rl4co_dihedral8_augment": ([_i64, _i32, _p, _p, _p], C.c_int),
    "rl4co_symmetric_augment": ([_i64, _i32, _p, _p, _p, _p], C.c_int),
    "rl4co_pomo_best": ([_i64, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p], C.c_int),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen the library and set signatures.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m rl4co_slap_b200.build` "
                "(there is no CPU fallback for the rollout path)")
        lib = C.CDLL(LIB_PATH)
        for name, (argtypes, restype) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = restype
        _lib = lib
    return _lib


def ptr(t: Optional[torch.Tensor], dtype: Optional[torch.dtype] = None) -> Optional[int]:
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("rl4co_slap_b200 kernels need CUDA tensors (no CPU fallback); got a "
                           f"{t.device} tensor")
    if not t.is_contiguous():
        raise RuntimeError("rl4co_slap_b200 kernels need contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected dtype {dtype}, got {t.dtype}")
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().rl4co_last_error().decode()
        raise RuntimeError(f"{what} failed (rc={rc}): {msg}")


def call(name: str, *args) -> None:
    check(getattr(load(), name)(*args), name)


def new_status(device) -> torch.Tensor:
    return torch.zeros(1, dtype=torch.int32, device=device)


_deferred = {}


def deferred_status(device) -> torch.Tensor:
    """Per-device status word the single-step env kernels OR into; it is read (one sync) by
    ``raise_deferred`` at the next get_reward, not after every step."""
    key = torch.device(device)
    if key.index is None:
        key = torch.device("cuda", torch.cuda.current_device())
    t = _deferred.get(key)
    if t is None:
        t = _deferred[key] = torch.zeros(1, dtype=torch.int32, device=key)
    return t


def raise_deferred(device) -> None:
    key = torch.device(device)
    if key.type != "cuda":
        return
    if key.index is None:
        key = torch.device("cuda", torch.cuda.current_device())
    t = _deferred.get(key)
    if t is not None:
        try:
            raise_for_status(t)
        finally:
            t.zero_()


def raise_for_status(status: torch.Tensor) -> None:
    """One device->host read per call; raises AssertionError with the reference's text."""
    bits = int(status.item())
    if bits:
        for bit, msg in STATUS_MESSAGES:
            if bits & bit:
                raise AssertionError(msg)
        raise AssertionError(f"device status {bits}")
