import itertools
from math import factorial


class MultiRoundShapleyValue:
    def __init__(self, players, last_round_metric=0.0, **kwargs) -> None:
        self.players = tuple(players)
        self.last_round_metric = float(last_round_metric)
        self.metric_fun = None
        self.shapley_values: dict = {}
        self.shapley_values_S: dict = {}

    def set_metric_function(self, fun) -> None:
        self.metric_fun = fun

    def compute(self, round_number: int) -> None:
        cache: dict = {}

        def metric(key: frozenset) -> float:
            if not key:
                return self.last_round_metric
            if key not in cache:
                cache[key] = float(self.metric_fun(tuple(sorted(key))))
            return cache[key]

        n = len(self.players)
        values = {p: 0.0 for p in self.players}
        for p in self.players:
            others = [q for q in self.players if q != p]
            for r in range(n):
                w = factorial(r) * factorial(n - r - 1) / factorial(n)
                for s in itertools.combinations(others, r):
                    values[p] += w * (metric(frozenset(s) | {p}) - metric(frozenset(s)))
        self.shapley_values = values
        best = max(cache, key=cache.get)
        self.shapley_values_S = {p: values[p] for p in sorted(best)}
        self.last_round_metric = metric(frozenset(self.players))
