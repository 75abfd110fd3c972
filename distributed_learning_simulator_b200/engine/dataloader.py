"""Device-resident batch iteration.

The whole phase tensor lives on the executor's device (shared between co-located
workers, see ``dataset._on_device``); an epoch is a device-side permutation plus
``index_select`` per batch, so iterating enqueues no host<->device copy and no
synchronisation.  ``storage="pinned_host"`` is the end-to-end mode: the worker's shard
sits in pinned host memory and is copied host->device at the start of every epoch.

Graph collections yield neighbour-sampled sub-graphs (``batch_number`` batches per epoch,
``num_neighbor`` expected in-neighbours per hop; knobs from
/root/reference/simulation_lib/worker/graph_worker.py:95-101).
"""
import math
from typing import Any, Iterator

import torch

from .dataset import DatasetCollection, GraphDatasetUtil
from .ml_type import MachineLearningPhase


class Batch:
    __slots__ = ("inputs", "kwargs", "targets", "batch_size", "output_rows")

    def __init__(self, inputs: tuple, targets: torch.Tensor, kwargs: dict | None = None,
                 output_rows: int | None = None) -> None:
        self.inputs = inputs
        self.kwargs = kwargs or {}
        self.targets = targets
        self.batch_size = int(targets.shape[0])
        self.output_rows = output_rows


class SampleLoader:
    def __init__(self, util: Any, batch_size: int, device: torch.device, shuffle: bool,
                 generator: torch.Generator | None, storage: str = "device") -> None:
        self.util = util
        self.batch_size = batch_size
        self.device = device
        self.shuffle = shuffle
        self.generator = generator
        self.storage = storage
        self.h2d_bytes = 0
        self._pinned: tuple[torch.Tensor, torch.Tensor] | None = None
        if storage == "pinned_host" and device.type == "cuda":
            cached = getattr(util, "_pinned_cache", None)
            if cached is None:
                cached = (util.inputs[util.indices].contiguous().pin_memory(),
                          util.targets[util.indices].contiguous().pin_memory())
                util._pinned_cache = cached
            self._pinned = cached

    def __len__(self) -> int:
        return math.ceil(len(self.util) / self.batch_size)

    @property
    def sample_count(self) -> int:
        return len(self.util)

    def epoch_tensors(self) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """(x_all, y_all, order): batches are ``x_all[order[i*B:(i+1)*B]]``."""
        if self._pinned is not None:
            stage = getattr(self.util, "_device_stage", None)
            if stage is None or stage[0].device != self.device:
                stage = (torch.empty_like(self._pinned[0], device=self.device),
                         torch.empty_like(self._pinned[1], device=self.device))
                self.util._device_stage = stage
            x, y = stage
            x.copy_(self._pinned[0], non_blocking=True)   # host -> device, every epoch
            y.copy_(self._pinned[1], non_blocking=True)
            self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
            n = y.shape[0]
            order = torch.arange(n, device=self.device)
        else:
            x, y, order = self.util.device_tensors(self.device)
            n = order.numel()
        if self.shuffle:
            perm = torch.randperm(n, device=self.device, generator=self.generator)
            order = order[perm]
        return x, y, order

    def __iter__(self) -> Iterator[Batch]:
        x, y, order = self.epoch_tensors()
        for start in range(0, order.numel(), self.batch_size):
            idx = order[start : start + self.batch_size]
            yield Batch((x.index_select(0, idx),), y.index_select(0, idx))


def sample_subgraph(edge_index: torch.Tensor, seeds: torch.Tensor, num_nodes: int, hops: int,
                    num_neighbor: int | None, generator: torch.Generator | None) -> tuple[torch.Tensor, torch.Tensor]:
    """k-hop in-neighbourhood of ``seeds``.  Returns (n_id, local edge_index) with the
    seeds occupying the first ``len(seeds)`` positions of ``n_id``."""
    device = edge_index.device
    in_set = torch.zeros(num_nodes, dtype=torch.bool, device=device)
    in_set[seeds] = True
    frontier = in_set.clone()
    edge_keep = torch.zeros(edge_index.shape[1], dtype=torch.bool, device=device)
    deg = None
    for _ in range(hops):
        cand = frontier[edge_index[1]]
        if num_neighbor is not None and num_neighbor > 0:
            if deg is None:
                deg = torch.zeros(num_nodes, device=device).index_add_(
                    0, edge_index[1], torch.ones(edge_index.shape[1], device=device))
            p = (num_neighbor / deg.clamp(min=1.0))[edge_index[1]]
            r = torch.rand(edge_index.shape[1], device=device, generator=generator)
            cand = cand & (r < p)
        edge_keep |= cand
        new_nodes = torch.zeros_like(in_set)
        new_nodes[edge_index[0, cand]] = True
        frontier = new_nodes & ~in_set
        in_set |= new_nodes
    rest = in_set.clone()
    rest[seeds] = False
    n_id = torch.cat([seeds, rest.nonzero(as_tuple=False).view(-1)])
    lookup = torch.full((num_nodes,), -1, dtype=torch.long, device=device)
    lookup[n_id] = torch.arange(n_id.numel(), device=device)
    sub = edge_index[:, edge_keep]
    return n_id, lookup[sub]


class GraphLoader:
    def __init__(self, util: GraphDatasetUtil, device: torch.device, shuffle: bool, hops: int,
                 generator: torch.Generator | None, batch_number: int | None = None,
                 batch_size: int | None = None, num_neighbor: int | None = None, full_graph: bool = False) -> None:
        self.util = util
        self.device = device
        self.shuffle = shuffle
        self.hops = hops
        self.generator = generator
        self.num_neighbor = num_neighbor
        self.full_graph = full_graph
        n = len(util)
        if batch_number is not None:
            self.batch_size = max(math.ceil(n / max(int(batch_number), 1)), 1)
        else:
            self.batch_size = int(batch_size or n) or 1
        self.h2d_bytes = 0

    def __len__(self) -> int:
        if self.full_graph:
            return 1
        return max(math.ceil(len(self.util) / self.batch_size), 1)

    @property
    def sample_count(self) -> int:
        return len(self.util)

    def __iter__(self) -> Iterator[Batch]:
        from .dataset import _on_device

        key = self.util._base_key
        x = _on_device(key, "graph.x", self.util.x, self.device)
        y = _on_device(key, "graph.y", self.util.y, self.device)
        edge_index = self.util.edge_index.to(self.device)
        nodes = self.util.indices.to(self.device)
        if self.full_graph:
            n_id = torch.arange(x.shape[0], device=self.device)
            yield Batch((x, edge_index), y[nodes], {"n_id": n_id, "output_index": nodes})
            return
        if self.shuffle:
            nodes = nodes[torch.randperm(nodes.numel(), device=self.device, generator=self.generator)]
        for start in range(0, nodes.numel(), self.batch_size):
            seeds = nodes[start : start + self.batch_size]
            n_id, local_ei = sample_subgraph(edge_index, seeds, x.shape[0], self.hops, self.num_neighbor,
                                             self.generator)
            yield Batch((x[n_id], local_ei), y[seeds], {"n_id": n_id}, output_rows=int(seeds.numel()))


def create_loader(dc: DatasetCollection, phase: MachineLearningPhase, batch_size: int, device: torch.device,
                  model: torch.nn.Module | None = None, generator: torch.Generator | None = None,
                  **kwargs: Any) -> Any:
    util = dc.get_dataset_util(phase)
    shuffle = phase == MachineLearningPhase.Training
    if dc.is_graph:
        hops = 2
        if model is not None:
            from ..models.graph import MessagePassing

            hops = max(sum(isinstance(m, MessagePassing) for m in model.modules()), 1)
        return GraphLoader(util, device, shuffle, hops, generator,
                           batch_number=kwargs.get("batch_number"),
                           batch_size=None if kwargs.get("batch_number") else batch_size,
                           num_neighbor=kwargs.get("num_neighbor"),
                           full_graph=not shuffle)
    return SampleLoader(util, batch_size, device, shuffle, generator, storage=kwargs.get("storage", "device"))
