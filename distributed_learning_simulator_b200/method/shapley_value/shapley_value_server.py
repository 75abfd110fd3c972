"""Servers for the Shapley methods: an ``AggregationServer`` that also evaluates the initial
model (``need_init_performance``, reference shapley_value_server.py:4-7) and owns the
matching algorithm (GTG_shapley_value_server.py:6-7, multiround_shapley_value_server.py:5-9)."""
from typing import Any

from ..common_import import AggregationServer
from .shapley_value_algorithm import (GTGShapleyValueAlgorithm, HierarchicalShapleyValueAlgorithm,
                                      MultiRoundShapleyValueAlgorithm)


class ShapleyValueServer(AggregationServer):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.need_init_performance = True


class GTGShapleyValueServer(ShapleyValueServer):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs, algorithm=GTGShapleyValueAlgorithm(server=self))


class MultiRoundShapleyValueServer(ShapleyValueServer):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs, algorithm=MultiRoundShapleyValueAlgorithm(server=self))


class HierarchicalShapleyValueServer(ShapleyValueServer):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs, algorithm=HierarchicalShapleyValueAlgorithm(server=self))
