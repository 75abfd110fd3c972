"""Shapley engines on toy games with known values."""
import itertools

import pytest

from distributed_learning_simulator_b200.algorithm.shapley_value import (GTGShapleyValue, HierarchicalShapleyValue,
                                                                          MultiRoundShapleyValue)


def additive(weights, base=0.0):
    return lambda subset: base + sum(weights[p] for p in subset)


def test_multiround_exact_on_additive_game():
    w = {0: 0.1, 1: 0.2, 2: 0.3, 3: 0.05}
    sv = MultiRoundShapleyValue(players=list(w), last_round_metric=0.0)
    sv.set_metric_function(additive(w))
    sv.compute(round_number=1)
    for p in w:
        assert sv.shapley_values[p] == pytest.approx(w[p])
    assert sum(sv.shapley_values.values()) == pytest.approx(sum(w.values()))  # efficiency
    assert sv.evaluations == 2 ** 4 - 1
    assert set(sv.shapley_values_S) == {0, 1, 2, 3}
    assert sv.last_round_metric == pytest.approx(sum(w.values()))


def test_multiround_symmetry_and_null_player():
    def game(subset):  # players 0 and 1 are substitutes, 2 is a null player
        s = set(subset)
        return 1.0 if (0 in s or 1 in s) else 0.0

    sv = MultiRoundShapleyValue(players=[0, 1, 2], last_round_metric=0.0)
    sv.set_metric_function(game)
    sv.compute(1)
    assert sv.shapley_values[0] == pytest.approx(0.5)
    assert sv.shapley_values[1] == pytest.approx(0.5)
    assert sv.shapley_values[2] == pytest.approx(0.0)


def test_gtg_matches_exact_on_small_game():
    w = {0: 0.10, 1: 0.25, 2: 0.05, 3: 0.15, 4: 0.20}

    def game(subset):  # additive plus a synergy between 0 and 1
        s = set(subset)
        return 0.1 + sum(w[p] for p in s) + (0.1 if {0, 1} <= s else 0.0)

    exact = MultiRoundShapleyValue(players=list(w), last_round_metric=0.1)
    exact.set_metric_function(game)
    exact.compute(1)
    gtg = GTGShapleyValue(players=list(w), last_round_metric=0.1, seed=3, eps=1e-9, converge_criteria=0.01)
    gtg.set_metric_function(game)
    gtg.compute(1)
    for p in w:
        assert gtg.shapley_values[p] == pytest.approx(exact.shapley_values[p], abs=0.03)
    assert sum(gtg.shapley_values.values()) == pytest.approx(game(w) - 0.1, abs=1e-9)
    # memoisation: never more evaluations than coalitions
    assert gtg.evaluations <= 2 ** 5 - 1


def test_gtg_between_round_truncation_and_batching():
    calls = []

    def game(subset):
        calls.append(tuple(subset))
        return 0.5

    gtg = GTGShapleyValue(players=[0, 1, 2], last_round_metric=0.5)
    gtg.set_metric_function(game)
    gtg.compute(1)
    assert all(v == 0.0 for v in gtg.shapley_values.values())
    batches = []

    def batch(subsets):
        batches.append(list(subsets))
        return [0.1 * len(s) for s in subsets]

    gtg2 = GTGShapleyValue(players=[0, 1, 2, 3], last_round_metric=0.0, seed=1)
    gtg2.set_metric_function(lambda s: 0.1 * len(s))
    gtg2.set_batch_metric_function(batch)
    gtg2.compute(1)
    assert len(batches[0]) == 5  # the grand coalition + the four singletons in one launch
    for p in range(4):
        assert gtg2.shapley_values[p] == pytest.approx(0.1)


def test_hierarchical_efficiency_and_groups():
    w = {i: 0.05 * (i + 1) for i in range(6)}
    h = HierarchicalShapleyValue(players=list(w), last_round_metric=0.0, part_number=3, vp_size=3)
    assert sorted(itertools.chain.from_iterable(h.groups)) == list(range(6)) and len(h.groups) == 3
    h.set_metric_function(additive(w))
    h.compute(1)
    for p in w:
        assert h.shapley_values[p] == pytest.approx(w[p])
    assert sum(h.shapley_values.values()) == pytest.approx(sum(w.values()))
    assert h.evaluations <= 2 ** 3 - 1 + 3 * (2 ** 2 - 1)
    h2 = HierarchicalShapleyValue(players=list(range(7)), part_number=2, vp_size=3)
    assert max(len(g) for g in h2.groups) <= 3


def test_gtg_speculative_prefetch_gives_the_sequential_values():
    """With a batched evaluator the GTG engine evaluates every prefix of a sweep up front and applies the
    truncation rule afterwards: Shapley values, best subset and the random permutation stream are those
    of the strictly sequential walk; only the number of utility evaluations grows."""
    from distributed_learning_simulator_b200.algorithm.shapley_value.gtg_shapley_value import GTGShapleyValue

    weights = {0: 0.30, 1: 0.05, 2: 0.22, 3: 0.0005, 4: 0.12, 5: 0.18}

    def utility(subset):
        s = sum(weights[p] for p in subset)
        return 0.1 + 0.8 * (1 - (1 - s) ** 2)  # concave: later members add less, so truncation kicks in

    results = {}
    for mode in ("sequential", "batched", "batched_no_speculation"):
        calls = {"single": 0, "batched": 0}

        def single(subset):
            calls["single"] += 1
            return utility(subset)

        def batch(subsets):
            calls["batched"] += len(subsets)
            return [utility(s) for s in subsets]

        engine = GTGShapleyValue(players=list(weights), last_round_metric=0.1, seed=7,
                                 speculative_prefetch=(mode == "batched"))
        engine.set_metric_function(single)
        engine.set_batch_metric_function(None if mode == "sequential" else batch)
        engine.compute(round_number=1)
        results[mode] = (dict(engine.shapley_values), dict(engine.shapley_values_S), dict(calls))
    seq, spec, plain = results["sequential"], results["batched"], results["batched_no_speculation"]
    assert seq[0] == spec[0] == plain[0] and seq[1] == spec[1] == plain[1]
    assert seq[2]["batched"] == 0 and spec[2]["single"] == 0 and plain[2]["single"] == 0
    assert spec[2]["batched"] >= plain[2]["batched"] >= 1
    assert abs(sum(seq[0].values()) - (utility(weights) - 0.1)) < 0.05  # efficiency, up to truncation
