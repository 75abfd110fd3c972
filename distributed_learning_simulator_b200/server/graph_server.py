"""``AggregationServer`` pre-bound to the node-embedding passing algorithm
(/root/reference/simulation_lib/server/graph_server.py:5-7)."""
from typing import Any

from ..algorithm.graph_algorithm import GraphNodeEmbeddingPassingAlgorithm
from .aggregation_server import AggregationServer


class GraphNodeServer(AggregationServer):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        kwargs.pop("algorithm", None)
        super().__init__(GraphNodeEmbeddingPassingAlgorithm(), *args, **kwargs)
