"""Shapley-value methods: FedAvg training whose server values every worker each round
(/root/reference/simulation_lib/method/shapley_value/__init__.py:6-15), plus
``Hierarchical_shapley_value`` which the reference's config names but never registers."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import AggregationWorker
from .shapley_value_algorithm import (GTGShapleyValueAlgorithm, HierarchicalShapleyValueAlgorithm,  # noqa: F401
                                      MultiRoundShapleyValueAlgorithm, ShapleyValueAlgorithm)
from .shapley_value_server import (GTGShapleyValueServer, HierarchicalShapleyValueServer,  # noqa: F401
                                   MultiRoundShapleyValueServer, ShapleyValueServer)

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="multiround_shapley_value", client_cls=AggregationWorker, server_cls=MultiRoundShapleyValueServer)
CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="GTG_shapley_value", client_cls=AggregationWorker, server_cls=GTGShapleyValueServer)
CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="Hierarchical_shapley_value", client_cls=AggregationWorker,
    server_cls=HierarchicalShapleyValueServer)
