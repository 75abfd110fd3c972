"""Single-model adaptive federated dropout (``single_model_afd``).  The reference keeps
configs (/root/reference/conf/smafd/*.yaml, conf/large_scale/smafd/*.yaml: ``dropout_rate``,
``random_client_number``) and two scaffolds -- ``ErrorFeedbackWorker``
(worker/error_feedback_worker.py:9-19) and ``RandomDropoutAlgorithm``
(algorithm/random_dropout_algorithm.py:7-31) -- but no registered method."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import AggregationServer, FedAVGAlgorithm
from .worker import SingleModelAFDWorker

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="single_model_afd", client_cls=SingleModelAFDWorker, server_cls=AggregationServer,
    algorithm_cls=FedAVGAlgorithm)
