"""Exact per-round Shapley values (multi-round SV): enumerate all 2^N - 1 non-empty
coalitions every round with ``U(empty) = last round's metric``."""
from .base import ShapleyValue


class MultiRoundShapleyValue(ShapleyValue):
    def compute(self, round_number: int) -> None:
        self.round_number = round_number
        self._cache = {}
        subsets = [frozenset(s) for s in self.powerset(self.players) if s]
        self.prefetch(subsets)
        self.shapley_values = self.exact_shapley(self.players, self.metric)
        best = self._best_subset(subsets)
        if best:
            members = tuple(sorted(best))
            self.shapley_values_S = self.exact_shapley(members, self.metric)
        else:
            self.shapley_values_S = {}
        self.last_round_metric = self.metric(frozenset(self.players))
