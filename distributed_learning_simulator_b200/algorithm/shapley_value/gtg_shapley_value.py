"""GTG-Shapley (Liu et al., ACM TIST 2022, doi 10.1145/3501811): guided truncated gradient
Shapley with between-round and within-round truncation.

Per round: ``v_N = U(all)``.  If ``|v_N - last_round_metric|`` is
below ``round_trunc_threshold`` every value is 0.  Otherwise sweeps of N guided
permutations (each player leads one permutation per sweep, the rest is a random shuffle
drawn from a seeded numpy RNG) are sampled; along a permutation ``v[j]`` is evaluated only
while ``|v_N - v[j-1]| >= eps`` (within-round truncation, later prefixes reuse ``v[j-1]``);
marginal contributions are recorded; sampling stops when the running means have converged
(last ``last_k`` means within ``converge_criteria`` relative change, at least
``converge_min`` records) or at the cap ``min(2^N, max(converge_min, 0.8 * 2^N))``.

Permutations led by different players are independent, so each sweep's first-level
prefixes -- and, sweep after sweep, all memoisation misses that are certain to be needed --
are handed to ``prefetch`` and evaluated in ONE batched launch when the server installed a
batched evaluator; the truncation rule then decides which values are used, giving the same
Shapley values as the strictly sequential evaluation of the reference
(method/shapley_value/shapley_value_algorithm.py:67-76).
"""
import numpy as np

from .base import ShapleyValue


class GTGShapleyValue(ShapleyValue):
    def __init__(self, players, last_round_metric: float = 0.0, eps: float = 0.001,
                 round_trunc_threshold: float = 0.001, converge_min: int | None = None, last_k: int = 10,
                 converge_criteria: float = 0.05, max_percentage: float = 0.8, seed: int = 0,
                 speculative_prefetch: bool = True, **kwargs) -> None:
        super().__init__(players, last_round_metric, **kwargs)
        self.speculative_prefetch = speculative_prefetch
        # every prefix of every surviving permutation is n^2 coalitions per sweep: worth it while a batched
        # launch is cheaper than the cache misses it avoids, not for hundreds of players
        self.speculative_max_players = 32
        self.eps = eps
        self.round_trunc_threshold = round_trunc_threshold
        self.converge_min = converge_min if converge_min is not None else max(30, self.player_number)
        self.last_k = last_k
        self.converge_criteria = converge_criteria
        self.max_percentage = max_percentage
        self.rng = np.random.RandomState(seed)
        self.max_number = min(2 ** self.player_number,
                              max(self.converge_min, int(self.max_percentage * 2 ** self.player_number)))

    def _converged(self, records: list[np.ndarray]) -> bool:
        if len(records) < self.converge_min:
            return False
        if len(records) >= self.max_number:
            return True
        arr = np.asarray(records)
        means = np.cumsum(arr, axis=0) / np.arange(1, len(arr) + 1).reshape(-1, 1)
        final = means[-1]
        tail = means[-self.last_k:]
        err = np.mean(np.abs(tail - final) / (np.abs(final) + 1e-12), axis=1)
        return bool(np.max(err) <= self.converge_criteria)

    def compute(self, round_number: int) -> None:
        self.round_number = round_number
        self._cache = {}
        players = list(self.players)
        n = len(players)
        self.prefetch([players] + [[p] for p in players])
        v_all = self.metric(players)
        if abs(v_all - self.last_round_metric) <= self.round_trunc_threshold:
            self.shapley_values = {p: 0.0 for p in players}
            self.shapley_values_S = {p: 0.0 for p in players}
            self.last_round_metric = v_all
            return
        records: list[np.ndarray] = []
        index = {p: i for i, p in enumerate(players)}
        while not self._converged(records):
            sweep = []
            for lead in players:
                rest = [p for p in players if p != lead]
                self.rng.shuffle(rest)
                sweep.append([lead] + rest)
            # level-2 prefixes of the whole sweep are needed unless truncated at level 1: batch them
            self.prefetch([perm[:2] for perm in sweep
                           if abs(v_all - self.metric(perm[:1])) >= self.eps and n >= 2])
            if self.batch_metric_fun is not None and self.speculative_prefetch and n <= self.speculative_max_players:
                # With a batched evaluator one more coalition costs a slice of a launch, while a cache miss
                # inside the walk below costs a launch of its own plus a host read: evaluate every prefix of
                # the sweep whose permutation survived level 2 up front.  The truncation rule still decides
                # which utilities are *used*, so the values are those of the sequential walk.
                alive = [perm for perm in sweep if n >= 3 and abs(v_all - self.metric(perm[:1])) >= self.eps
                         and abs(v_all - self.metric(perm[:2])) >= self.eps]
                self.prefetch([perm[:j] for perm in alive for j in range(3, n + 1)])
            for perm in sweep:
                marginal = np.zeros(n)
                prev = self.last_round_metric
                for j in range(1, n + 1):
                    cur = self.metric(perm[:j]) if abs(v_all - prev) >= self.eps else prev
                    marginal[index[perm[j - 1]]] = cur - prev
                    prev = cur
                records.append(marginal)
                if self._converged(records):
                    break
        sv = np.mean(np.asarray(records), axis=0)
        self.shapley_values = {p: float(sv[index[p]]) for p in players}
        best = self._best_subset(self._cache.keys())
        if best:
            members = tuple(sorted(best))
            if len(members) <= 6:
                self.prefetch([c for c in self.powerset(members) if c])
            # restrict to the best coalition; unseen sub-coalitions fall back to evaluation
            self.shapley_values_S = self.exact_shapley(members, self.metric) if len(members) <= 6 else {
                p: self.shapley_values[p] for p in members}
        else:
            self.shapley_values_S = {}
        self.last_round_metric = v_all
