"""Hierarchical (two-level) Shapley values for ``Hierarchical_shapley_value``
(config: /root/reference/conf/hierarchical_sv/mnist.yaml:4,13-15 -- ``part_number: 3``,
``vp_size: 3``, 6 workers; the reference has no implementation).

Players are dealt into ``part_number`` virtual players of at most ``vp_size`` members.
Level 1: exact Shapley values of the virtual players, where a coalition of virtual players
is the union of their members.  Level 2: inside each virtual player, exact Shapley values
of its members on the sub-game ``U(S)`` for ``S`` within the group, rescaled so that the
members' values sum to the group's level-1 value (efficiency is preserved at both levels).
Cost: ``2^part_number + sum_g 2^|g|`` utilities instead of ``2^N``.
"""
import math

from .base import ShapleyValue


class HierarchicalShapleyValue(ShapleyValue):
    def __init__(self, players, last_round_metric: float = 0.0, part_number: int = 2, vp_size: int | None = None,
                 **kwargs) -> None:
        super().__init__(players, last_round_metric, **kwargs)
        n = self.player_number
        part_number = max(1, min(int(part_number), n))
        if vp_size is not None and part_number * int(vp_size) < n:
            part_number = math.ceil(n / int(vp_size))
        self.groups: list[tuple] = [tuple(self.players[i::part_number]) for i in range(part_number)]
        self.group_values: dict = {}

    def compute(self, round_number: int) -> None:
        self.round_number = round_number
        self._cache = {}
        group_ids = list(range(len(self.groups)))

        def union(coalition) -> frozenset:
            return frozenset(p for g in coalition for p in self.groups[g])

        needed = [union(c) for c in self.powerset(group_ids) if c]
        for members in self.groups:
            needed += [frozenset(s) for s in self.powerset(members) if s]
        self.prefetch(needed)
        self.group_values = self.exact_shapley(group_ids, lambda c: self.metric(union(c)))
        values: dict = {}
        for g, members in enumerate(self.groups):
            inner = self.exact_shapley(members, self.metric)
            total = sum(inner.values())
            target = self.group_values[g]
            if abs(total) > 1e-12:
                values.update({p: v * target / total for p, v in inner.items()})
            else:
                values.update({p: target / len(members) for p in members})
        self.shapley_values = values
        best = self._best_subset(self._cache.keys())
        self.shapley_values_S = {p: values[p] for p in sorted(best)} if best else {}
        self.last_round_metric = self.metric(frozenset(self.players))
