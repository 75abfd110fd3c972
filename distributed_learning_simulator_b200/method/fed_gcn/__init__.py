"""FedGCN (convergence/communication trade-offs in federated GCN training): the graph worker
with feature sharing forced on (/root/reference/simulation_lib/method/fed_gcn/__init__.py:8-12,
worker.py:4-7)."""
from typing import Any

from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import GraphNodeServer, GraphWorker


class FedGCNWorker(GraphWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._share_feature = True


CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_gcn", client_cls=FedGCNWorker, server_cls=GraphNodeServer)
