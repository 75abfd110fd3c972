"""FedPAQ (periodic averaging + quantization, arXiv:1909.13014): FedAvg whose uploads and
downloads pass through 255-level stochastic-quantization endpoints
(/root/reference/simulation_lib/method/fed_paq/__init__.py:7-14)."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import (AggregationServer, AggregationWorker, FedAVGAlgorithm, StochasticQuantClientEndpoint,
                             StochasticQuantServerEndpoint)

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_paq", client_cls=AggregationWorker, server_cls=AggregationServer,
    client_endpoint_cls=StochasticQuantClientEndpoint, server_endpoint_cls=StochasticQuantServerEndpoint,
    algorithm_cls=FedAVGAlgorithm)
