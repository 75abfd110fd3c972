"""fed_aas: sub-graph training with in-client edge dropping and optional embedding sharing.
The reference ships configs (/root/reference/conf/fed_aas/*.yaml: ``share_feature``,
``batch_number``, ``edge_drop_rate``, ``num_neighbor``) but registers nothing; every knob is
implemented by the graph worker (worker/graph_worker.py:22,95-101,205-216), so the method is
that worker with sharing OFF unless the config asks for it."""
from typing import Any

from ..algorithm_factory import CentralizedAlgorithmFactory
from ..common_import import GraphNodeServer, GraphWorker


class FedAASWorker(GraphWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._share_feature = self.config.algorithm_kwargs.get("share_feature", False)


CentralizedAlgorithmFactory.register_algorithm(
    algorithm_name="fed_aas", client_cls=FedAASWorker, server_cls=GraphNodeServer)
