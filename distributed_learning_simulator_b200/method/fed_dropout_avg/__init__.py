"""FedDropoutAvg (arXiv:2111.13230): uploads are randomly masked element-wise and the
server averages every element over the workers whose upload kept it
(/root/reference/simulation_lib/method/fed_dropout_avg/__init__.py:7-12)."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import AggregationServer
from .algorithm import FedDropoutAvgAlgorithm
from .worker import FedDropoutAvgWorker

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_dropout_avg", client_cls=FedDropoutAvgWorker, server_cls=AggregationServer,
    algorithm_cls=FedDropoutAvgAlgorithm)
