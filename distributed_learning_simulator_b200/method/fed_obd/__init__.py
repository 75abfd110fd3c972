"""FedOBD (arXiv:2208.05174): opportunistic block dropout + quantized exchange, two phases.
``fed_obd`` uses NNADQ endpoints, ``fed_obd_sq`` stochastic quantization
(/root/reference/simulation_lib/method/fed_obd/__init__.py:8-22)."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import (NNADQClientEndpoint, NNADQServerEndpoint, StochasticQuantClientEndpoint,
                             StochasticQuantServerEndpoint)
from .obd_algorithm import OpportunisticBlockDropoutAlgorithm, get_module_blocks  # noqa: F401
from .phase import Phase  # noqa: F401
from .server import FedOBDServer
from .worker import FedOBDWorker

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_obd", client_cls=FedOBDWorker, server_cls=FedOBDServer,
    client_endpoint_cls=NNADQClientEndpoint, server_endpoint_cls=NNADQServerEndpoint)
CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_obd_sq", client_cls=FedOBDWorker, server_cls=FedOBDServer,
    client_endpoint_cls=StochasticQuantClientEndpoint, server_endpoint_cls=StochasticQuantServerEndpoint)
