"""fed_gnn = (GraphWorker, GraphNodeServer)
(/root/reference/simulation_lib/method/fed_gnn/__init__.py:4-8)."""
from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import GraphNodeServer, GraphWorker

CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_gnn", client_cls=GraphWorker, server_cls=GraphNodeServer)
