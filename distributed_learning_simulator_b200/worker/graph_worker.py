"""Node-partitioned federated GNN training over one global graph.

Behaviour parity with /root/reference/simulation_lib/worker/graph_worker.py:18-406:
exchange of training-node index sets through the server (:68-84); pruning of the training
graph to in-client edges (optionally Bernoulli-dropped with ``edge_drop_rate``), cross-client
edges and validation edges (:197-241); without ``share_feature`` the cross-client edges are
removed statically (:243-250); with it, layer 0 strips them on the fly (:252-269) and every
later message-passing layer swaps boundary-node embeddings through the server each batch
(:344-373); byte accounting and ``graph_worker_stat.json`` (:382-406).

Differences by design: node sets are boolean masks / index tensors on the trainer's device,
so gathering boundary rows and merging received embeddings are ``index_select`` /
``index_copy`` instead of Python loops over ``n_id.tolist()`` with a host sync per layer per
batch (:271-321); the node-mask size is the number of nodes, not the number of edges
(SURVEY.md defect D6); a config without ``batch_number`` trains full-batch instead of
raising ``KeyError`` (:95-97).
"""
import functools
import json
import os
from typing import Any, Callable

import torch

from ..ops import graph as graph_ops
from ..engine.dataset import GraphDatasetUtil, coalesce, index_to_mask, mask_to_index
from ..engine.log import get_logger
from ..engine.ml_type import MachineLearningPhase
from ..message import Message, ParameterMessageBase, get_message_size
from ..models.graph import MessagePassing
from .aggregation_worker import AggregationWorker


class GraphWorker(AggregationWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if self.config.trainer_config.hook_config.use_amp:
            get_logger().warning("graph workers train in fp32; use_amp is ignored")
            self.config.trainer_config.hook_config.use_amp = False
        self._share_feature = self.config.algorithm_kwargs.get("share_feature", True)
        self._other_training_node_indices: set = set()
        self._other_training_node_mask: torch.Tensor | None = None
        self.__old_edge_index: None | torch.Tensor = None
        self.__n_id: None | torch.Tensor = None
        self._hook_handles: dict = {}
        self._comunicated_batch_cnt: int = 0
        self._skipped_embedding_bytes: int = 0
        self._round_skipped_bytes: dict = {}
        self._communicated_embedding_bytes: int = 0
        self._aggregated_bytes: int = 0
        self._round_communicated_bytes: dict = {}
        self.__local_node_mask: None | torch.Tensor = None
        self.__boundary_mask: None | torch.Tensor = None
        self._original_in_client_training_edge_cnt: int = 0
        self._in_client_training_edge_cnt: int = 0
        self._cross_client_training_edge_cnt: int = 0
        self._recorded_model_size: dict = {}
        self.__prepared = False

    # ---- dataset views ---------------------------------------------------------------------
    def get_dataset_util(self, phase: MachineLearningPhase) -> GraphDatasetUtil:
        util = self.trainer.dataset_collection.get_dataset_util(phase=phase)
        assert isinstance(util, GraphDatasetUtil)
        return util

    @property
    def num_nodes(self) -> int:
        return self.get_dataset_util(MachineLearningPhase.Training).num_nodes

    @property
    def edge_index(self) -> torch.Tensor:
        return self.get_dataset_util(phase=MachineLearningPhase.Training).get_edge_index(graph_index=0)

    @functools.cached_property
    def training_node_mask(self) -> torch.Tensor:
        return self.get_dataset_util(phase=MachineLearningPhase.Training).get_mask()[0]

    @functools.cached_property
    def validation_node_mask(self) -> torch.Tensor | None:
        if not self.trainer.dataset_collection.has_dataset(phase=MachineLearningPhase.Validation):
            return None
        return self.get_dataset_util(phase=MachineLearningPhase.Validation).get_mask()[0]

    @functools.cached_property
    def training_node_indices(self) -> set:
        return set(mask_to_index(self.training_node_mask).tolist())

    @property
    def local_node_mask(self) -> torch.Tensor:
        if self.__local_node_mask is None:
            mask = self.training_node_mask.clone()
            if self.validation_node_mask is not None:
                mask |= self.validation_node_mask
            self.__local_node_mask = mask
        return self.__local_node_mask

    @property
    def other_training_node_mask(self) -> torch.Tensor:
        assert self._other_training_node_mask is not None
        return self._other_training_node_mask

    @property
    def training_boundary_mask(self) -> torch.Tensor:
        """My training nodes that have an edge into another client's training nodes."""
        if self.__boundary_mask is None:
            ei = self.edge_index
            cross = self.training_node_mask[ei[0]] & self.other_training_node_mask[ei[1]]
            self.__boundary_mask = index_to_mask(ei[0, cross], self.num_nodes)
        return self.__boundary_mask

    @property
    def training_node_boundary(self) -> set:
        return set(mask_to_index(self.training_boundary_mask).tolist())

    @property
    def in_client_edge_mask(self) -> torch.Tensor:
        ei = self.edge_index
        return self.training_node_mask[ei[0]] & self.training_node_mask[ei[1]]

    @property
    def cross_client_edge_mask(self) -> torch.Tensor:
        ei = self.edge_index
        return self.training_node_mask[ei[0]] & self.other_training_node_mask[ei[1]]

    # ---- set-up --------------------------------------------------------------------------------
    def __exchange_training_node_indices(self) -> None:
        self.send_data_to_server(Message(other_data={"training_node_indices": self.training_node_indices},
                                         in_round=True))
        res = self._get_data_from_server()
        assert isinstance(res, Message)
        for worker_id, node_indices in res.other_data["training_node_indices"].items():
            if worker_id != self.worker_id:
                self._other_training_node_indices |= set(node_indices)
        assert not self._other_training_node_indices.intersection(self.training_node_indices)
        other = torch.tensor(sorted(self._other_training_node_indices), dtype=torch.long)
        self._other_training_node_mask = index_to_mask(other, self.num_nodes)

    def _before_training(self) -> None:
        super()._before_training()
        if self.__prepared or self._stopped():
            return
        self.__prepared = True
        if self.worker_id == 0:
            get_logger().warning("share feature" if self._share_feature else "not share feature")
        self.__exchange_training_node_indices()
        self.__clear_unrelated_edges()
        kwargs = self.config.algorithm_kwargs
        self.trainer.update_dataloader_kwargs(batch_number=kwargs.get("batch_number", 1))
        if "num_neighbor" in kwargs:
            self.trainer.update_dataloader_kwargs(num_neighbor=kwargs["num_neighbor"])
        for idx, module in enumerate(self._get_message_passing_modules()):
            module.register_forward_pre_hook(
                hook=functools.partial(self._record_embedding_size, module_index=idx), with_kwargs=True)
        if not self._share_feature:
            self._clear_cross_client_edges()
            return
        self.trainer.model.register_forward_pre_hook(hook=self._catch_n_id, with_kwargs=True, prepend=True)
        for idx, _ in enumerate(self._get_message_passing_modules()):
            hook: Callable = self._pass_node_feature
            if idx == 0:
                hook = self._clear_cross_client_edge_on_the_fly
            self._register_embedding_hook(module_index=idx, hook=hook)

    def _register_embedding_hook(self, module_index: int, hook: Callable) -> None:
        old_hook = self._hook_handles.pop(module_index, None)
        if old_hook is not None:
            old_hook.remove()
        self._hook_handles[module_index] = self._get_message_passing_modules()[module_index].register_forward_pre_hook(
            hook=functools.partial(hook, module_index=module_index), with_kwargs=True)

    def __replace_training_edges(self, edge_index: torch.Tensor) -> None:
        self.trainer.dataset_collection.transform_dataset(
            self.trainer.phase,
            lambda _, dataset_util, __: dataset_util.get_edge_subset(graph_index=0, edge_index=edge_index))

    def __clear_unrelated_edges(self) -> None:
        # Keep in-device edges and cross-device edges
        edge_index = self.edge_index
        in_client = self.in_client_edge_mask
        cross_client = self.cross_client_edge_mask
        self._original_in_client_training_edge_cnt = int(in_client.sum().item())
        new_in_client = in_client.clone()
        edge_drop_rate: float | None = self.config.algorithm_kwargs.get("edge_drop_rate", None)
        if edge_drop_rate is not None and edge_drop_rate != 0:
            if self.worker_id == 0:
                get_logger().warning("drop in client edge with rate: %s", edge_drop_rate)
            g = torch.Generator().manual_seed(self.config.reproducible_env_config.seed * 7919 + self.worker_id)
            new_in_client &= torch.rand(new_in_client.shape, generator=g) < (1 - edge_drop_rate)
        if new_in_client.sum().item() != 0:
            get_logger().warning("cross_client_edge/in_client_edge %s",
                                 cross_client.sum().item() / new_in_client.sum().item())
        self._cross_client_training_edge_cnt = int(cross_client.sum().item())
        self._in_client_training_edge_cnt = int(new_in_client.sum().item())
        edge_mask = new_in_client | cross_client
        if self.validation_node_mask is not None:
            edge_mask |= self.validation_node_mask[edge_index[0]] & self.validation_node_mask[edge_index[1]]
        self.__boundary_mask = None
        _ = self.training_boundary_mask  # computed on the unpruned graph, like the reference
        self.__replace_training_edges(coalesce(edge_index[:, edge_mask], self.num_nodes))

    def _clear_cross_client_edges(self) -> None:
        edge_index = self.edge_index
        local = self.local_node_mask
        keep = local[edge_index[0]] & local[edge_index[1]]
        self.__replace_training_edges(coalesce(edge_index[:, keep], self.num_nodes))

    # ---- forward hooks ------------------------------------------------------------------------------
    @property
    def n_id(self) -> torch.Tensor:
        assert self.__n_id is not None
        return self.__n_id

    def _catch_n_id(self, module: torch.nn.Module, args: Any, kwargs: Any) -> tuple | None:
        self.__n_id = kwargs["n_id"]
        return args, kwargs

    def _on_device(self, mask: torch.Tensor, device: torch.device) -> torch.Tensor:
        cache = self.__dict__.setdefault("_mask_cache", {})
        key = (id(mask), str(device))
        if key not in cache:
            cache[key] = mask.to(device)
        return cache[key]

    def _clear_cross_client_edge_on_the_fly(self, module: torch.nn.Module, args: Any, kwargs: Any,
                                            module_index: int) -> tuple | None:
        if not module.training:
            return None
        self.__old_edge_index = args[1]
        n_id = self.n_id
        ei = args[1]
        # K13: endpoints are looked up through n_id and the kept edges compacted in order on the device
        kept = graph_ops.filter_edges(ei, self._on_device(self.local_node_mask, n_id.device), n_id)
        if module_index > 0:
            x = args[0]
            cnt = int(self._on_device(self.training_boundary_mask, n_id.device)[n_id].sum().item())
            self._skipped_embedding_bytes += x.shape[1] * x.element_size() * cnt
        return (args[0], kept, *args[2:]), kwargs

    def training_boundary_feature(self, x: torch.Tensor) -> tuple | None:
        n_id = self.n_id
        assert x.shape[0] == n_id.shape[0]
        # K13 node select + K14 row gather
        rows, ids = graph_ops.select_nodes(n_id, self._on_device(self.training_boundary_mask, n_id.device))
        if rows.numel() == 0:
            return None
        return graph_ops.gather_rows(x, rows), ids

    def _get_cross_device_embedding(self, embedding_indices: Any, embedding: torch.Tensor | None,
                                    x: torch.Tensor) -> torch.Tensor:
        n_id = self.n_id
        train_mask = self._on_device(self.training_node_mask, n_id.device)
        ids = None
        if embedding is not None and len(embedding_indices) != 0:
            ids = torch.as_tensor(embedding_indices, dtype=torch.long, device=n_id.device)
            embedding = embedding.to(x.device, non_blocking=True)
        # K14: foreign rows take the received embedding of that node (zero when nobody sent one)
        return graph_ops.halo_merge(x, n_id, train_mask, ids, embedding if ids is not None else None)

    def _record_embedding_size(self, module: torch.nn.Module, args: Any, kwargs: Any,
                               module_index: int) -> tuple | None:
        if "model_bytes" not in self._recorded_model_size:
            parameter = self.trainer.model_util.get_parameter_list()
            self._recorded_model_size["model_bytes"] = parameter.element_size() * parameter.numel()
        sizes = self._recorded_model_size.setdefault("embedding_bytes", {})
        if module_index not in sizes:
            x = args[0]
            sizes[module_index] = x[0].numel() * x[0].element_size()
        return None

    def _pass_node_feature(self, module: torch.nn.Module, args: Any, kwargs: Any, module_index: int) -> tuple | None:
        if not module.training:
            return None
        x = args[0]
        n_id = self.n_id
        wanted = n_id[self._on_device(self.other_training_node_mask, n_id.device)[n_id]]
        feature = self.training_boundary_feature(x)
        sent_data = Message(other_data={"node_embedding": feature, "boundary": wanted}, in_round=True)
        cnt = 0 if feature is None else int(feature[1].numel())
        self._comunicated_batch_cnt += 1
        self._communicated_embedding_bytes += cnt * x.shape[1] * x.element_size()
        self.send_data_to_server(sent_data)
        res = self._get_data_from_server()
        assert res is not None
        new_x = self._get_cross_device_embedding(res["node_indices"], res["node_embedding"], x)
        edge_index = self.__old_edge_index if self.__old_edge_index is not None else args[1]
        return (new_x, edge_index, *args[2:]), kwargs

    def _get_message_passing_modules(self) -> list:
        return [module for module in self.trainer.model.modules() if isinstance(module, MessagePassing)]

    # ---- accounting ------------------------------------------------------------------------------------
    def _get_sent_data(self) -> ParameterMessageBase:
        sent_data = super()._get_sent_data()
        self._aggregated_bytes += get_message_size(sent_data)
        self._round_communicated_bytes[self._round_num] = self._aggregated_bytes + self._communicated_embedding_bytes
        self._round_skipped_bytes[self._round_num] = self._skipped_embedding_bytes
        return sent_data

    def _after_training(self) -> None:
        super()._after_training()
        stat = {
            "original_in_client_training_edge_cnt": self._original_in_client_training_edge_cnt,
            "in_client_training_edge_cnt": self._in_client_training_edge_cnt,
            "cross_client_training_edge_cnt": self._cross_client_training_edge_cnt,
            "training_node_cnt": int(self.training_node_mask.sum().item()),
            "skipped_embedding_bytes": self._round_skipped_bytes,
            "communicated_bytes": self._round_communicated_bytes,
        }
        if self.validation_node_mask is not None:
            stat["validation_node_cnt"] = int(self.validation_node_mask.sum().item())
        with open(os.path.join(self.save_dir, "graph_worker_stat.json"), "wt", encoding="utf8") as f:
            json.dump(stat | self._recorded_model_size, f)
