"""rl4co_am_rollout_static: the rollout with its table scratch in a static __device__ array of the library
(no counterpart in the reference; added for callers that cannot allocate).

Called through the C ABI (ctypes) and compared with the existing entry points rl4co_am_project_tc +
rl4co_am_rollout on caller-owned tables, on two calls in a row with different inputs (the array keeps the first
call's contents).  Same kernels and launchers behind both, so tours are compared exactly and log-likelihoods /
rewards at the suite's 1e-5 relative tolerance.  The file sorts last on purpose.
"""

import pytest
import torch

import rl4co_slap_b200 as R
from rl4co_slap_b200 import _lib
from tests.helpers import load_policy

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _args(pol, env, td, h, mode, S, seed, tables, out):
    """AmRolloutArgs for a tsp / cvrp rollout, filled like AttentionModelPolicy.forward does."""
    name, (B, N, _) = env.name, td["locs"].shape
    W, q0, wcap, _ = pol.decoder.folded_weights()
    a = _lib.AmRolloutArgs()
    a.env, a.mode, a.B, a.S, a.multistart = _lib.ENV_IDS[name], mode, B, S, int(S > 0)
    a.num_loc = N - (0 if name == "tsp" else 1)
    a.N, a.T, a.P, a.O, a.K, a.n_tables = N, out["actions"].shape[1], 0, 0, 0, W.shape[0] // 128
    a.temperature, a.tanh_clipping, a.mask_logits = 1.0, 10.0, 1
    a.tables, a.gctx = _lib.ptr(tables), _lib.ptr(out["gctx"])
    a.q0, a.wcap = _lib.ptr(q0), _lib.ptr(wcap)
    a.locs = _lib.ptr(td["locs"].contiguous(), torch.float32)
    if name == "cvrp":
        a.demand = _lib.ptr(td["demand"].contiguous(), torch.float32)
        a.vehicle_capacity = _lib.ptr(td["vehicle_capacity"].contiguous(), torch.float32)
    a.seed, a.offset = seed, 0
    a.actions, a.logp, a.ll = _lib.ptr(out["actions"]), _lib.ptr(out["logp"]), _lib.ptr(out["ll"])
    a.reward, a.steps, a.status = _lib.ptr(out["reward"]), _lib.ptr(out["steps"]), _lib.ptr(out["status"])
    return a


def _outputs(pol, h, R_, T, dev):
    B, N = h.shape[0], h.shape[1]
    out = {"actions": torch.full((R_, T), -1, dtype=torch.int64, device=dev),
           "logp": torch.zeros(R_, T, device=dev), "ll": torch.zeros(R_, device=dev),
           "reward": torch.zeros(R_, device=dev), "steps": torch.zeros(R_, dtype=torch.int32, device=dev),
           "status": _lib.new_status(dev), "gctx": None}
    if pol.decoder.use_graph_context:
        out["gctx"] = torch.empty(B, 128, device=dev)
        _lib.call("rl4co_am_graph_context", B, N, _lib.ptr(h), _lib.ptr(pol.decoder.folded_weights()[3]),
                  _lib.ptr(out["gctx"]), _lib.stream_ptr(dev))
    return out


def _both(name, B, N, mode, S, seed, dev, gen_seed):
    """-> (existing entry points' outputs, static entry point's outputs) on the same fresh inputs."""
    torch.manual_seed(gen_seed)
    pol = load_policy(name, dev)
    env = R.get_env(name, device=dev, generator_params=dict(num_loc=N - (0 if name == "tsp" else 1)))
    td = env.reset(batch_size=[B])
    with torch.inference_mode():
        h = pol.encoder(td)[0].float().contiguous()
    image, stream = pol.decoder.weight_image(), _lib.stream_ptr(dev)
    n_tables = pol.decoder.folded_weights()[0].shape[0] // 128
    T = N if name == "tsp" else 2 * (N - 1) + 1
    rows = B * max(S, 1)
    # existing path: caller-owned tables
    ref = _outputs(pol, h, rows, T, dev)
    tables = torch.empty(n_tables, B * N, 128, device=dev)
    _lib.call("rl4co_am_project_tc", B * N, n_tables, _lib.ptr(h), _lib.ptr(image), _lib.ptr(tables), stream)
    a = _args(pol, env, td, h, mode, S, seed, tables, ref)
    _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
    # new path: no tables pointer, scratch inside the library
    got = _outputs(pol, h, rows, T, dev)
    b = _args(pol, env, td, h, mode, S, seed, None, got)
    _lib.check(_lib.load().rl4co_am_rollout_static(b, _lib.ptr(h), _lib.ptr(image), stream), "rl4co_am_rollout_static")
    torch.cuda.synchronize()
    return ref, got


def _compare(ref, got):
    assert int(ref["status"].item()) == 0 and int(got["status"].item()) == 0
    assert torch.equal(ref["steps"], got["steps"])
    assert torch.equal(ref["actions"], got["actions"]), "tours differ between the static-scratch and the regular entry point"
    assert (got["actions"] >= 0).all()
    torch.testing.assert_close(got["ll"], ref["ll"], rtol=RTOL, atol=1e-5)
    torch.testing.assert_close(got["reward"], ref["reward"], rtol=RTOL, atol=1e-6)
    torch.testing.assert_close(got["logp"], ref["logp"], rtol=RTOL, atol=2e-5)


def test_static_scratch_rollout_matches_regular_entry_point_twice(dev):
    # call 1: TSP50 greedy, 100 instances (12.8 MB of the 32 MiB array)
    _compare(*_both("tsp", 100, 50, _lib.MODE_GREEDY, 0, 0, dev, gen_seed=21))
    # call 2, straight after, other env / sizes / mode: CVRP20 POMO-style multistart sampling (same Philox seed on both
    # sides -> identical draws); the array still holds call 1's tables where call 2 does not overwrite them
    _compare(*_both("cvrp", 37, 21, _lib.MODE_SAMPLING, 8, 1234, dev, gen_seed=22))
    # and the first configuration again with new instances: stale contents of call 2 must not leak in
    _compare(*_both("tsp", 64, 50, _lib.MODE_GREEDY, 0, 0, dev, gen_seed=23))


def test_static_scratch_rejects_oversize_and_caller_tables(dev):
    torch.manual_seed(3)
    pol = load_policy("tsp", dev)
    env = R.TSPEnv(generator_params=dict(num_loc=100), device=dev)
    B, N = 132, 100                                     # 5 tables x 132 x 100 x 512 B = 33 792 000 B > 32 MiB >= 131 instances
    assert 5 * B * N * 512 > _lib.STATIC_TABLE_BYTES >= 5 * (B - 1) * N * 512
    td = env.reset(batch_size=[B])
    h = torch.zeros(B, N, 128, device=dev)
    out = _outputs(pol, h, B, N, dev)
    stream, image = _lib.stream_ptr(dev), pol.decoder.weight_image()
    a = _args(pol, env, td, h, _lib.MODE_GREEDY, 0, 0, None, out)
    rc = _lib.load().rl4co_am_rollout_static(a, _lib.ptr(h), _lib.ptr(image), stream)
    assert rc != 0
    msg = _lib.load().rl4co_last_error().decode()
    assert "RL4CO_STATIC_TABLE_BYTES" in msg and str(_lib.STATIC_TABLE_BYTES) in msg, msg
    assert (out["actions"] == -1).all(), "a rejected call must not launch anything"
    # a caller-supplied tables pointer is an error too (that is what rl4co_am_rollout is for)
    small = torch.empty(5, 4 * N, 128, device=dev)
    a = _args(pol, env, td[:4], h[:4], _lib.MODE_GREEDY, 0, 0, small, out)
    assert _lib.load().rl4co_am_rollout_static(a, _lib.ptr(h), _lib.ptr(image), stream) != 0
    assert "must be NULL" in _lib.load().rl4co_last_error().decode()
    # the largest request that fits runs
    a = _args(pol, env, td[:B - 1], h[:B - 1], _lib.MODE_GREEDY, 0, 0, None, out)
    _lib.check(_lib.load().rl4co_am_rollout_static(a, _lib.ptr(h), _lib.ptr(image), stream), "rl4co_am_rollout_static")
    torch.cuda.synchronize()
    assert int(out["status"].item()) == 0
    assert torch.equal(out["actions"][:B - 1].sort(1).values, torch.arange(N, device=dev).expand(B - 1, N))
