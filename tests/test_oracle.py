"""CPU: oracle/port.py against the committed fixtures generated from the unmodified
reference (tests/golden, oracle/gen_golden.py).  This is what pins the oracle where
/root/reference is absent."""
import numpy as np
import pytest
import torch

from oracle import port
from tests.helpers import (ATSP_CASES, CVRP_CASES, OP_CASES, SDVRP_CASES, PCTSP_CASES, SLAP_CASES, TSP_CASES, env_of, fixture, oracle_reset,
                           oracle_weights, t)


def _state_keys(fx, prefix):
    return [k[len(prefix) + 1:] for k in fx.files if k.startswith(prefix + ".")]


@pytest.mark.parametrize("case", TSP_CASES + CVRP_CASES + SLAP_CASES + OP_CASES + PCTSP_CASES + ATSP_CASES + SDVRP_CASES)
def test_replay_bit_exact(case):
    """every per-step state tensor (mask, visited, capacity, i, done ...) equals the reference's"""
    fx, name = fixture(case), env_of(case)
    td = oracle_reset(name, fx)
    for k in _state_keys(fx, "reset"):
        if k in td:
            assert np.array_equal(td[k].numpy(), fx[f"reset.{k}"]), (case, "reset", k)
    acts = t(fx["replay.actions"])
    for step in range(int(fx["replay.steps"])):
        td["action"] = acts[:, step]
        td = port._STEP[name](td)
        for k in _state_keys(fx, f"step{step}"):
            if k in td:
                got = td[k].numpy()
                assert got.dtype == fx[f"step{step}.{k}"].dtype, (case, step, k)
                assert np.array_equal(got, fx[f"step{step}.{k}"]), (case, step, k)
    if name == "tsp":
        r = port.tsp_reward(td["locs"], acts)
    elif name == "cvrp":
        port.cvrp_check(td["demand"], td["vehicle_capacity"], acts)
        r = port.cvrp_reward(td["locs"], acts)
    elif name == "op":
        port.op_check(td["locs"], td["max_length"], acts)
        r = port.op_reward(td["prize"], acts)
    elif name in ("pctsp", "spctsp"):
        port.pctsp_check(td, acts)
        r = port.pctsp_reward(td, acts)
    elif name == "atsp":
        port.atsp_check(acts)
        r = port.atsp_reward(td["cost_matrix"], acts)
    elif name == "sdvrp":
        port.sdvrp_check(td["demand"], td["vehicle_capacity"], acts)
        r = port.cvrp_reward(td["locs"], acts)
    else:
        r = port.slap_reward(td["locs"], td["assignment"], td["picklist"])
    assert np.array_equal(r.numpy(), fx["replay.reward"])


@pytest.mark.parametrize("case", ("tsp20", "tsp50", "tsp100", "cvrp20", "cvrp50", "slap50", "slap100") + OP_CASES + PCTSP_CASES + SDVRP_CASES)
def test_greedy_rollout_matches_reference(case):
    fx, name = fixture(case), env_of(case)
    out = port.am_rollout(name, oracle_reset(name, fx), t(fx["hidden"]), oracle_weights(name), "greedy")
    assert np.array_equal(out["actions"].numpy(), fx["greedy.actions"])
    np.testing.assert_allclose(out["reward"].numpy(), fx["greedy.reward"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(out["logprobs"].numpy(), fx["greedy.logprobs"], rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("case", ("tsp20", "cvrp20", "op20", "op50", "op100", "pctsp20", "pctsp100", "spctsp50"))
def test_multistart_greedy_matches_reference(case):
    fx, name = fixture(case), env_of(case)
    S = int(fx["ms_greedy.num_starts"])
    out = port.am_rollout(name, oracle_reset(name, fx), t(fx["hidden"]), oracle_weights(name),
                          "multistart_greedy", num_starts=S)
    assert np.array_equal(out["actions"].numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["reward"].numpy(), fx["ms_greedy.reward"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(out["log_likelihood"].numpy(), fx["ms_greedy.ll"], rtol=1e-5, atol=1e-5)


def test_pomo_policy_multistart():
    fx = fixture("cvrp20_pomo")
    out = port.am_rollout("cvrp", oracle_reset("cvrp", fx), t(fx["hidden"]),
                          oracle_weights("cvrp", "cvrp_pomo"), "multistart_greedy", num_starts=20,
                          use_graph_context=False)
    assert np.array_equal(out["actions"].numpy(), fx["ms_greedy.actions"])
    np.testing.assert_allclose(out["log_likelihood"].numpy(), fx["ms_greedy.ll"], rtol=1e-5, atol=1e-5)


def test_evaluate_reproduces_ll():
    fx = fixture("tsp20")
    out = port.am_rollout("tsp", oracle_reset("tsp", fx), t(fx["hidden"]), oracle_weights("tsp"),
                          actions=t(fx["greedy.actions"]))
    np.testing.assert_allclose(out["log_likelihood"].numpy(), fx["greedy.ll"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("case", OP_CASES + PCTSP_CASES + SDVRP_CASES)
def test_op_evaluate_long_tours(case):
    """per-step log-probs of the replay's random tours (they run into the length budget) = the reference's"""
    fx, name = fixture(case), env_of(case)
    out = port.am_rollout(name, oracle_reset(name, fx), t(fx["hidden"]), oracle_weights(name),
                          actions=t(fx["replay.actions"]))
    np.testing.assert_allclose(out["logprobs"].numpy(), fx["eval.logprobs"], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(out["reward"].numpy(), fx["eval.reward"], rtol=1e-6, atol=1e-6)


def test_invalid_tour_raises():
    fx = fixture("tsp5")
    bad = t(fx["replay.actions"]).clone()
    bad[0, 1] = bad[0, 0]
    with pytest.raises(AssertionError, match="Invalid tour"):
        port.tsp_reward(t(fx["gen.locs"]), bad)
