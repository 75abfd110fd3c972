"""fed_avg = (AggregationWorker, AggregationServer, FedAVGAlgorithm)
(/root/reference/simulation_lib/method/fed_avg/__init__.py:5-10)."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import AggregationServer, AggregationWorker, FedAVGAlgorithm

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_avg", client_cls=AggregationWorker, server_cls=AggregationServer,
    algorithm_cls=FedAVGAlgorithm)
