"""Sequential GTG-Shapley (same rules as SURVEY.md appendix D), one utility call at a time --
the way the reference evaluates coalitions."""
import numpy as np


class GTGShapleyValue:
    def __init__(self, players, last_round_metric=0.0, eps=0.001, round_trunc_threshold=0.001, last_k=10,
                 converge_criteria=0.05, max_percentage=0.8, seed=0, **kwargs) -> None:
        self.players = tuple(players)
        self.last_round_metric = float(last_round_metric)
        self.eps, self.round_trunc_threshold = eps, round_trunc_threshold
        self.last_k, self.converge_criteria = last_k, converge_criteria
        n = len(self.players)
        self.converge_min = max(30, n)
        self.max_number = min(2 ** n, max(self.converge_min, int(max_percentage * 2 ** n)))
        self.rng = np.random.RandomState(seed)
        self.metric_fun = None
        self.shapley_values: dict = {}
        self.shapley_values_S: dict = {}

    def set_metric_function(self, fun) -> None:
        self.metric_fun = fun

    def _converged(self, records) -> bool:
        if len(records) < self.converge_min:
            return False
        if len(records) >= self.max_number:
            return True
        arr = np.asarray(records)
        means = np.cumsum(arr, 0) / np.arange(1, len(arr) + 1).reshape(-1, 1)
        err = np.mean(np.abs(means[-self.last_k:] - means[-1]) / (np.abs(means[-1]) + 1e-12), axis=1)
        return bool(np.max(err) <= self.converge_criteria)

    def compute(self, round_number: int) -> None:
        cache: dict = {}

        def metric(subset) -> float:
            key = frozenset(subset)
            if not key:
                return self.last_round_metric
            if key not in cache:
                cache[key] = float(self.metric_fun(tuple(sorted(key))))
            return cache[key]

        players = list(self.players)
        n = len(players)
        v_all = metric(players)
        if abs(v_all - self.last_round_metric) <= self.round_trunc_threshold:
            self.shapley_values = {p: 0.0 for p in players}
            self.shapley_values_S = dict(self.shapley_values)
            self.last_round_metric = v_all
            return
        index = {p: i for i, p in enumerate(players)}
        records: list = []
        while not self._converged(records):
            for lead in players:
                rest = [p for p in players if p != lead]
                self.rng.shuffle(rest)
                perm = [lead] + rest
                marginal, prev = np.zeros(n), self.last_round_metric
                for j in range(1, n + 1):
                    cur = metric(perm[:j]) if abs(v_all - prev) >= self.eps else prev
                    marginal[index[perm[j - 1]]] = cur - prev
                    prev = cur
                records.append(marginal)
                if self._converged(records):
                    break
        sv = np.mean(np.asarray(records), 0)
        self.shapley_values = {p: float(sv[index[p]]) for p in players}
        best = max(cache, key=cache.get)
        self.shapley_values_S = {p: self.shapley_values[p] for p in sorted(best)}
        self.last_round_metric = v_all
