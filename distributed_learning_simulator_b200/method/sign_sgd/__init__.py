"""signSGD with majority vote (Bernstein et al., 2018).  The reference ships configs
(/root/reference/conf/sign_sgd/*.yaml: one "round" of 100 epochs, no initial broadcast)
and the per-step ``GradientWorker`` scaffold but registers no implementation
(SURVEY.md 2.4); this package provides it."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from .server import SignSGDServer
from .worker import SignSGDWorker

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="sign_SGD", client_cls=SignSGDWorker, server_cls=SignSGDServer)
