import itertools
from typing import Callable, Iterable, Sequence


class ShapleyValue:
    def __init__(self, players: Iterable, last_round_metric: float = 0.0, **kwargs) -> None:
        self.players: tuple = tuple(players)
        self.last_round_metric: float = float(last_round_metric)
        self.metric_fun: Callable | None = None
        self.batch_metric_fun: Callable | None = None  # evaluates many subsets in one launch
        self.shapley_values: dict = {}
        self.shapley_values_S: dict = {}
        self.round_number = 0
        self._cache: dict[frozenset, float] = {}
        self.evaluations = 0

    @property
    def player_number(self) -> int:
        return len(self.players)

    def set_metric_function(self, metric_fun: Callable) -> None:
        self.metric_fun = metric_fun

    def set_batch_metric_function(self, batch_metric_fun: Callable | None) -> None:
        self.batch_metric_fun = batch_metric_fun

    def metric(self, subset: Iterable) -> float:
        key = frozenset(subset)
        if not key:
            return self.last_round_metric
        if key not in self._cache:
            if self.batch_metric_fun is not None:  # same evaluator (and rounding) as the prefetched values
                self._cache[key] = float(self.batch_metric_fun([tuple(sorted(key))])[0])
            else:
                assert self.metric_fun is not None
                self._cache[key] = float(self.metric_fun(tuple(sorted(key))))
            self.evaluations += 1
        return self._cache[key]

    def prefetch(self, subsets: Sequence[Iterable]) -> None:
        """Evaluate the not-yet-known subsets together when a batched evaluator is present."""
        missing = []
        for s in subsets:
            key = frozenset(s)
            if key and key not in self._cache and key not in missing:
                missing.append(key)
        if not missing:
            return
        if self.batch_metric_fun is not None:
            values = self.batch_metric_fun([tuple(sorted(k)) for k in missing])
            for k, v in zip(missing, values):
                self._cache[k] = float(v)
            self.evaluations += len(missing)
        else:
            for k in missing:
                self.metric(k)

    def compute(self, round_number: int) -> None:
        raise NotImplementedError

    @staticmethod
    def powerset(players: Sequence) -> Iterable[tuple]:
        return itertools.chain.from_iterable(itertools.combinations(players, r) for r in range(len(players) + 1))

    def exact_shapley(self, players: Sequence, metric: Callable[[frozenset], float]) -> dict:
        """Classic weighted marginal sum over all coalitions of ``players``."""
        from math import factorial

        n = len(players)
        values = {p: 0.0 for p in players}
        for p in players:
            others = [q for q in players if q != p]
            for r in range(n):
                w = factorial(r) * factorial(n - r - 1) / factorial(n)
                for s in itertools.combinations(others, r):
                    values[p] += w * (metric(frozenset(s) | {p}) - metric(frozenset(s)))
        return values

    def _best_subset(self, candidates: Iterable[frozenset]) -> frozenset:
        best, best_v = frozenset(), float("-inf")
        for s in candidates:
            if not s:
                continue
            v = self._cache.get(frozenset(s))
            if v is not None and v > best_v:
                best, best_v = frozenset(s), v
        return best
