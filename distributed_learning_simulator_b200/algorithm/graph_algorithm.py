"""Server-side node-embedding exchange for federated GNN training.

Behaviour of /root/reference/simulation_lib/algorithm/graph_algorithm.py:7-89: three
message kinds share the aggregation slot -- (a) ``training_node_indices`` exchange,
(b) per-batch boundary-embedding exchange answered with a per-worker unicast under
``other_data["worker_result"]``, (c) ordinary FedAvg of parameters.  The reference
indexes embeddings row by row in Python dicts and stacks them on the CPU
(graph_algorithm.py:36-42,74-81); here all uploaded rows are concatenated once and each
worker's answer is one device-side gather through a global-id -> row lookup table.
"""
import torch

from ..message import Message, ParameterMessageBase
from .fed_avg_algorithm import FedAVGAlgorithm


class GraphNodeEmbeddingPassingAlgorithm(FedAVGAlgorithm):
    def __init__(self) -> None:
        super().__init__()
        self.__embeddings: list[torch.Tensor] = []
        self.__embedding_ids: list[torch.Tensor] = []
        self.__boundaries: dict[int, set] = {}

    def process_worker_data(self, worker_id: int, worker_data: Message | None, old_parameter_dict: dict | None,
                            save_dir: str) -> None:
        super().process_worker_data(worker_id=worker_id, worker_data=worker_data,
                                    old_parameter_dict=old_parameter_dict, save_dir=save_dir)
        data = self._all_worker_data.get(worker_id)
        if data is None:
            return
        if "node_embedding" in data.other_data:
            node_embedding_tuple = data.other_data.pop("node_embedding")
            if node_embedding_tuple is not None:
                node_embedding, node_indices = node_embedding_tuple
                ids = torch.as_tensor(node_indices, dtype=torch.long, device=node_embedding.device)
                self.__embeddings.append(node_embedding)
                self.__embedding_ids.append(ids)
            self.__boundaries[worker_id] = data.other_data.pop("boundary")

    def aggregate_worker_data(self) -> Message:
        first = next(iter(self._all_worker_data.values()))
        if not isinstance(first, ParameterMessageBase):
            if "training_node_indices" in first.other_data:
                return Message(in_round=True, other_data={"training_node_indices": {
                    worker_id: data.other_data["training_node_indices"]
                    for worker_id, data in self._all_worker_data.items()}})
            res: dict = {"worker_result": {}}
            if self.__embeddings:
                device = self.__embeddings[0].device
                table = torch.cat([e.to(device) for e in self.__embeddings])
                ids = torch.cat([i.to(device) for i in self.__embedding_ids])
                order = torch.argsort(ids)
                sorted_ids = ids[order]
            for worker_id, boundary in self.__boundaries.items():
                node_indices: tuple = ()
                node_embedding = None
                if self.__embeddings and len(boundary):
                    if isinstance(boundary, torch.Tensor):
                        want = torch.unique(boundary.to(sorted_ids.device))
                    else:
                        want = torch.as_tensor(sorted(boundary), dtype=torch.long, device=sorted_ids.device)
                    pos = torch.searchsorted(sorted_ids, want).clamp(max=sorted_ids.numel() - 1)
                    hit = sorted_ids[pos] == want
                    if bool(hit.any()):
                        node_indices = tuple(want[hit].tolist())
                        node_embedding = table[order[pos[hit]]]
                res["worker_result"][worker_id] = {"node_indices": node_indices, "node_embedding": node_embedding}
            for worker_id in self._all_worker_data:
                res["worker_result"].setdefault(worker_id, {"node_indices": (), "node_embedding": None})
            self.__embeddings, self.__embedding_ids, self.__boundaries = [], [], {}
            return Message(in_round=True, other_data=res)
        return super().aggregate_worker_data()
