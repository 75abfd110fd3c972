"""Synthetic dataset collections with the shapes of the datasets the reference names.

The reference gets its datasets from the external ``cyy_torch_vision/text/graph``
packages (/root/reference/simulation_lib/method/common_import.py:1-2,13-16) and
downloads MNIST/CIFAR/IMDB/Planetoid at run time.  There is no network here, and the
benchmark contract is "synthetic data of the named dataset shape", so every dataset is
generated deterministically from a seed: class-conditional prototypes plus noise (the
models can actually learn them, which lets the tests assert that accuracy rises).

A collection keeps whole-phase tensors (``inputs``/``targets``) plus an index subset per
phase.  Base tensors are cached per (dataset, kwargs) and per device, so a hundred
simulated workers packed on one B200 share ONE copy of the data in HBM and differ only
in their index tensors.

API surface mirrored from ``cyy_torch_toolbox.dataset`` as consumed by the reference:
``name``, ``get_labels``, ``is_mutilabel`` (sic), ``has_dataset``, ``remove_dataset``,
``transform_dataset``, ``get_dataset_util`` (config.py:57-64, practitioner.py:32-34,
sampler/base.py:16-38, worker/aggregation_worker.py:27-34, worker/graph_worker.py:39-47).
"""
import copy
import math
import threading
from typing import Any, Callable

import torch

from .ml_type import MachineLearningPhase, ModelType

_cache_lock = threading.Lock()
_base_cache: dict[tuple, dict] = {}
_device_cache: dict[tuple, torch.Tensor] = {}


def _freeze(obj: Any) -> Any:
    if isinstance(obj, dict):
        return tuple(sorted((k, _freeze(v)) for k, v in obj.items()))
    if isinstance(obj, (list, tuple)):
        return tuple(_freeze(v) for v in obj)
    return obj


# name -> (kind, spec).  Sizes are the real datasets' sizes unless noted.
DATASET_SPECS: dict[str, dict] = {
    "MNIST": dict(kind="vision", shape=(1, 28, 28), classes=10, train=60000, test=10000),
    "CIFAR10": dict(kind="vision", shape=(3, 32, 32), classes=10, train=50000, test=10000),
    "CIFAR100": dict(kind="vision", shape=(3, 32, 32), classes=100, train=50000, test=10000),
    # ImageNet-shape; default sample count is reduced (1.28 M x 3x224x224 fp32 = 770 GB)
    "ImageNet": dict(kind="vision", shape=(3, 224, 224), classes=1000, train=12800, test=2000),
    "imdb": dict(kind="text", max_len=300, vocab=20000, classes=2, train=25000, test=25000),
    "IMDB": dict(kind="text", max_len=300, vocab=20000, classes=2, train=25000, test=25000),
    "Cora": dict(kind="graph", nodes=2708, features=1433, classes=7, avg_degree=3.9),
    "PubMed": dict(kind="graph", nodes=19717, features=500, classes=3, avg_degree=4.5),
    "Coauthor_CS": dict(kind="graph", nodes=18333, features=6805, classes=15, avg_degree=8.9),
    # the four below are scaled down from the real graphs (0.2-1.6 M nodes) by default
    "CitationFull": dict(kind="graph", nodes=17716, features=1639, classes=4, avg_degree=6.0),
    "Reddit": dict(kind="graph", nodes=23296, features=602, classes=41, avg_degree=20.0),
    "Yelp": dict(kind="graph", nodes=35842, features=300, classes=100, avg_degree=10.0),
    "AmazonProducts": dict(kind="graph", nodes=39261, features=200, classes=107, avg_degree=10.0),
}


def register_dataset(name: str, **spec: Any) -> None:
    DATASET_SPECS[name] = spec


def _generator(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def _make_vision(spec: dict, seed: int) -> dict:
    c, h, w = spec["shape"]
    classes = spec["classes"]
    g = _generator(seed)
    # low-frequency class prototypes, upsampled: learnable by small CNNs in a few steps
    lo = torch.randn(classes, c, max(h // 4, 1), max(w // 4, 1), generator=g)
    proto = torch.nn.functional.interpolate(lo, size=(h, w), mode="bilinear", align_corners=False)
    proto = proto / proto.flatten(1).std(dim=1).view(-1, 1, 1, 1)
    signal = float(spec.get("signal", 0.6))
    out = {}
    for phase, n in (("train", spec["train"]), ("test", spec["test"])):
        y = torch.randint(0, classes, (n,), generator=g)
        x = torch.randn(n, c, h, w, generator=g)
        x.add_(proto[y], alpha=signal)
        out[phase] = (x, y)
    return out


def _make_text(spec: dict, seed: int) -> dict:
    g = _generator(seed)
    vocab, max_len, classes = spec["vocab"], spec["max_len"], spec["classes"]
    # every class owns a band of "signal" tokens that appear with elevated probability
    band = max(vocab // (4 * classes), 1)
    out = {}
    for phase, n in (("train", spec["train"]), ("test", spec["test"])):
        y = torch.randint(0, classes, (n,), generator=g)
        tokens = torch.randint(2, vocab, (n, max_len), generator=g)
        sig = torch.rand(n, max_len, generator=g) < float(spec.get("signal", 0.15))
        sig_tok = 2 + y.view(-1, 1) * band + torch.randint(0, band, (n, max_len), generator=g)
        tokens = torch.where(sig, sig_tok, tokens)
        # variable lengths, 0 = padding
        lengths = torch.randint(max_len // 3, max_len + 1, (n, 1), generator=g)
        pos = torch.arange(max_len).view(1, -1)
        tokens = torch.where(pos < lengths, tokens, torch.zeros_like(tokens))
        out[phase] = (tokens, y)
    return out


def _make_graph(spec: dict, seed: int) -> dict:
    g = _generator(seed)
    n, f, classes = int(spec["nodes"]), int(spec["features"]), int(spec["classes"])
    y = torch.randint(0, classes, (n,), generator=g)
    proto = torch.randn(classes, f, generator=g)
    x = torch.randn(n, f, generator=g) + float(spec.get("signal", 0.5)) * proto[y]
    # stochastic-block-model edges: 80% intra-class
    e_half = int(n * float(spec["avg_degree"]) / 2)
    src = torch.randint(0, n, (e_half,), generator=g)
    intra = torch.rand(e_half, generator=g) < 0.8
    order = torch.argsort(y, stable=True)
    counts = torch.bincount(y, minlength=classes)
    starts = torch.cumsum(counts, 0) - counts
    r = torch.rand(e_half, generator=g)
    ys = y[src]
    same = order[(starts[ys] + (r * counts[ys].clamp(min=1)).long()).clamp(max=n - 1)]
    other = torch.randint(0, n, (e_half,), generator=g)
    dst = torch.where(intra, same, other)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    edge_index = torch.stack([torch.cat([src, dst]), torch.cat([dst, src])])
    edge_index = coalesce(edge_index, n)
    perm = torch.randperm(n, generator=g)
    n_train, n_val = int(0.6 * n), int(0.2 * n)
    masks = {}
    for phase, idx in (
        (MachineLearningPhase.Training, perm[:n_train]),
        (MachineLearningPhase.Validation, perm[n_train : n_train + n_val]),
        (MachineLearningPhase.Test, perm[n_train + n_val :]),
    ):
        m = torch.zeros(n, dtype=torch.bool)
        m[idx] = True
        masks[phase] = m
    return {"x": x, "y": y, "edge_index": edge_index, "masks": masks}


def coalesce(edge_index: torch.Tensor, num_nodes: int | None = None) -> torch.Tensor:
    """Sort edges by (row, col) and drop duplicates (own ``torch_geometric.utils.coalesce``)."""
    if edge_index.numel() == 0:
        return edge_index
    if num_nodes is None:
        num_nodes = int(edge_index.max().item()) + 1
    key = edge_index[0] * num_nodes + edge_index[1]
    key = torch.unique(key, sorted=True)
    return torch.stack([key // num_nodes, key % num_nodes])


def mask_to_index(mask: torch.Tensor) -> torch.Tensor:
    return mask.nonzero(as_tuple=False).view(-1)


def index_to_mask(index: torch.Tensor, size: int) -> torch.Tensor:
    mask = torch.zeros(size, dtype=torch.bool, device=index.device)
    mask[index] = True
    return mask


def _load_base(name: str, kwargs: dict) -> dict:
    if name not in DATASET_SPECS:
        raise KeyError(f"unknown dataset {name}; known: {sorted(DATASET_SPECS)}")
    key = (name, _freeze(kwargs))
    with _cache_lock:
        if key in _base_cache:
            return _base_cache[key]
        spec = dict(DATASET_SPECS[name])
        for k in ("train", "test", "nodes", "features", "classes", "vocab", "max_len", "signal", "avg_degree"):
            alias = {"train": "train_size", "test": "test_size", "nodes": "num_nodes"}.get(k, k)
            if alias in kwargs:
                spec[k] = kwargs[alias]
        seed = int(kwargs.get("seed", 20240229)) + sum(map(ord, name))
        if spec["kind"] == "vision":
            base = _make_vision(spec, seed)
        elif spec["kind"] == "text":
            base = _make_text(spec, seed)
        else:
            base = _make_graph(spec, seed)
        base["spec"] = spec
        base["key"] = key
        _base_cache[key] = base
        return base


def clear_dataset_cache() -> None:
    with _cache_lock:
        _base_cache.clear()
        _device_cache.clear()


def _on_device(base_key: tuple, tag: str, tensor: torch.Tensor, device: torch.device) -> torch.Tensor:
    """One resident copy of a base tensor per device, shared by all workers there."""
    if tensor.device == device:
        return tensor
    key = (base_key, tag, str(device))
    with _cache_lock:
        cached = _device_cache.get(key)
        if cached is None:
            cached = tensor.to(device, non_blocking=True)
            _device_cache[key] = cached
        return cached


class DatasetUtil:
    """A phase of a sample-indexed dataset: base tensors + an index subset."""

    def __init__(self, inputs: torch.Tensor, targets: torch.Tensor, indices: torch.Tensor | None,
                 base_key: tuple, tag: str) -> None:
        self.inputs = inputs
        self.targets = targets
        self.indices = indices if indices is not None else torch.arange(len(targets))
        self._base_key = base_key
        self._tag = tag

    def __len__(self) -> int:
        return int(self.indices.numel())

    def get_labels(self) -> set[int]:
        return set(torch.unique(self.targets[self.indices]).tolist())

    def get_label_indices(self) -> dict[int, torch.Tensor]:
        t = self.targets[self.indices]
        return {int(l): self.indices[t == l] for l in torch.unique(t).tolist()}

    def subset(self, indices: torch.Tensor) -> "DatasetUtil":
        return DatasetUtil(self.inputs, self.targets, indices, self._base_key, self._tag)

    def device_tensors(self, device: torch.device) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        x = _on_device(self._base_key, self._tag + ".x", self.inputs, device)
        y = _on_device(self._base_key, self._tag + ".y", self.targets, device)
        return x, y, self.indices.to(device)


class GraphDatasetUtil:
    """A phase of a single-graph node-classification dataset (``cyy_torch_graph``'s
    ``GraphDatasetUtil`` contract: get_edge_index / get_mask / get_edge_subset,
    /root/reference/simulation_lib/worker/graph_worker.py:39-47,157,176)."""

    def __init__(self, x: torch.Tensor, y: torch.Tensor, edge_index: torch.Tensor, mask: torch.Tensor,
                 base_key: tuple) -> None:
        self.x, self.y, self.edge_index, self.mask = x, y, edge_index, mask
        self._base_key = base_key

    def __len__(self) -> int:
        return int(self.mask.sum().item())

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    def get_labels(self) -> set[int]:
        return set(torch.unique(self.y).tolist())

    def get_label_indices(self) -> dict[int, torch.Tensor]:
        idx = mask_to_index(self.mask)
        t = self.y[idx]
        return {int(l): idx[t == l] for l in torch.unique(t).tolist()}

    def get_edge_index(self, graph_index: int = 0) -> torch.Tensor:
        return self.edge_index

    def get_mask(self) -> list[torch.Tensor]:
        return [self.mask]

    def get_edge_subset(self, graph_index: int, edge_index: torch.Tensor) -> "GraphDatasetUtil":
        return GraphDatasetUtil(self.x, self.y, edge_index, self.mask, self._base_key)

    def subset(self, indices: torch.Tensor) -> "GraphDatasetUtil":
        return GraphDatasetUtil(self.x, self.y, self.edge_index, index_to_mask(indices, self.num_nodes),
                                self._base_key)

    @property
    def indices(self) -> torch.Tensor:
        return mask_to_index(self.mask)


class DatasetCollection:
    def __init__(self, name: str, datasets: dict, spec: dict, dataset_kwargs: dict | None = None) -> None:
        self.name = name
        self._datasets: dict[MachineLearningPhase, Any] = datasets
        self.spec = spec
        self.dataset_kwargs = dataset_kwargs or {}

    @property
    def dataset_type(self) -> ModelType:
        return {
            "vision": ModelType.Classification,
            "text": ModelType.TextClassification,
            "graph": ModelType.GraphNodeClassification,
        }[self.spec["kind"]]

    @property
    def is_graph(self) -> bool:
        return self.spec["kind"] == "graph"

    def has_dataset(self, phase: MachineLearningPhase) -> bool:
        return phase in self._datasets

    def remove_dataset(self, phase: MachineLearningPhase) -> None:
        self._datasets.pop(phase, None)

    def get_dataset_util(self, phase: MachineLearningPhase = MachineLearningPhase.Test) -> Any:
        return self._datasets[phase]

    def set_dataset_util(self, phase: MachineLearningPhase, util: Any) -> None:
        self._datasets[phase] = util

    def transform_dataset(self, phase: MachineLearningPhase, transformer: Callable) -> None:
        """``transformer(dataset, dataset_util, phase) -> new dataset util``
        (call shape from worker/graph_worker.py:236-241)."""
        util = self._datasets[phase]
        self._datasets[phase] = transformer(util, util, phase)

    def get_dataset_size(self, phase: MachineLearningPhase) -> int:
        return len(self._datasets[phase])

    def get_labels(self) -> set[int]:
        return set(range(int(self.spec["classes"])))

    @property
    def label_number(self) -> int:
        return int(self.spec["classes"])

    def is_mutilabel(self) -> bool:  # spelling follows the reference (sampler/base.py:17)
        return False

    def set_subset(self, phase: MachineLearningPhase, indices: torch.Tensor) -> None:
        self._datasets[phase] = self._datasets[phase].subset(indices)

    def merge_validation_to_training(self) -> None:
        if not self.has_dataset(MachineLearningPhase.Validation):
            return
        tr = self._datasets[MachineLearningPhase.Training]
        va = self._datasets.pop(MachineLearningPhase.Validation)
        if self.is_graph:
            self._datasets[MachineLearningPhase.Training] = GraphDatasetUtil(
                tr.x, tr.y, tr.edge_index, tr.mask | va.mask, tr._base_key)
        else:
            raise NotImplementedError("merge is only defined for graph collections here")

    def clone(self) -> "DatasetCollection":
        return DatasetCollection(self.name, dict(self._datasets), self.spec, copy.deepcopy(self.dataset_kwargs))


class ClassificationDatasetCollection(DatasetCollection):
    pass


def create_dataset_collection(name: str, dataset_kwargs: dict | None = None) -> ClassificationDatasetCollection:
    dataset_kwargs = dict(dataset_kwargs or {})
    gen_kwargs = {k: v for k, v in dataset_kwargs.items() if k not in ("tokenizer", "name")}
    base = _load_base(name, gen_kwargs)
    spec, key = base["spec"], base["key"]
    datasets: dict = {}
    if spec["kind"] == "graph":
        for phase, mask in base["masks"].items():
            datasets[phase] = GraphDatasetUtil(base["x"], base["y"], base["edge_index"], mask, key)
    else:
        x, y = base["train"]
        datasets[MachineLearningPhase.Training] = DatasetUtil(x, y, None, key, "train")
        x, y = base["test"]
        # like the toolbox the reference builds on, datasets without an official
        # validation split donate half of their test split to validation
        n = len(y)
        half = n // 2
        datasets[MachineLearningPhase.Validation] = DatasetUtil(x, y, torch.arange(0, half), key, "test")
        datasets[MachineLearningPhase.Test] = DatasetUtil(x, y, torch.arange(half, n), key, "test")
    return ClassificationDatasetCollection(dataset_kwargs.get("name", name), datasets, spec, dataset_kwargs)


# ---------------------------------------------------------------------------------
# samplers (``cyy_torch_toolbox.dataset`` sampler registry; reference call sites
# config.py:58-63, practitioner.py:32-34, sampler/base.py:9-46)
# ---------------------------------------------------------------------------------
class _PhaseSampler:
    def __init__(self, util: Any, seed: int) -> None:
        self._util = util
        self._generator = _generator(seed)

    def iid_split_indices(self, part_number: int) -> list[torch.Tensor]:
        return self.split_indices([{l: 1 for l in self._util.get_labels()} for _ in range(part_number)])

    def random_split_indices(self, part_number: int) -> list[torch.Tensor]:
        idx = self._util.indices
        perm = idx[torch.randperm(idx.numel(), generator=self._generator)]
        return [t.clone() for t in torch.tensor_split(perm, part_number)]

    def split_indices(self, parts: list[dict]) -> list[torch.Tensor]:
        """``parts[i]`` maps label -> weight; each label's samples are divided between the
        parts that list it, proportionally to the weights."""
        label_indices = self._util.get_label_indices()
        result: list[list[torch.Tensor]] = [[] for _ in parts]
        for label, idx in sorted(label_indices.items()):
            owners = [(i, p[label]) for i, p in enumerate(parts) if p.get(label, 0) > 0]
            if not owners:
                continue
            idx = idx[torch.randperm(idx.numel(), generator=self._generator)]
            total_w = float(sum(w for _, w in owners))
            start = 0
            for j, (i, w) in enumerate(owners):
                end = idx.numel() if j == len(owners) - 1 else start + int(math.floor(idx.numel() * w / total_w))
                result[i].append(idx[start:end])
                start = end
        return [torch.sort(torch.cat(r))[0] if r else torch.empty(0, dtype=torch.long) for r in result]


class DatasetCollectionSampler:
    def __init__(self, dataset_collection: DatasetCollection, part_number: int, seed: int = 0, **kwargs: Any) -> None:
        self._dc: DatasetCollection | None = None
        self._part_number = part_number
        self._dataset_indices: dict[MachineLearningPhase, dict[int, torch.Tensor]] = {}
        self._samplers: dict[MachineLearningPhase, _PhaseSampler] = {}
        self._seed = seed
        self.set_dataset_collection(dataset_collection)

    def set_dataset_collection(self, dataset_collection: DatasetCollection) -> None:
        self._dc = dataset_collection
        if not self._samplers:
            for i, phase in enumerate(MachineLearningPhase):
                if dataset_collection.has_dataset(phase):
                    self._samplers[phase] = _PhaseSampler(dataset_collection.get_dataset_util(phase), self._seed + i)

    def __getstate__(self) -> dict:
        state = self.__dict__.copy()
        state["_dc"] = None
        return state

    def sample(self, part_id: int) -> None:
        assert self._dc is not None
        for phase, parts in self._dataset_indices.items():
            if self._dc.has_dataset(phase):
                self._dc.set_subset(phase, parts[part_id])


class IIDSampler(DatasetCollectionSampler):
    """Label-stratified equal split of every phase."""

    def __init__(self, dataset_collection: DatasetCollection, part_number: int, **kwargs: Any) -> None:
        super().__init__(dataset_collection, part_number, **kwargs)
        for phase, s in self._samplers.items():
            self._dataset_indices[phase] = dict(enumerate(s.iid_split_indices(part_number)))


class RandomSampler(DatasetCollectionSampler):
    def __init__(self, dataset_collection: DatasetCollection, part_number: int, **kwargs: Any) -> None:
        super().__init__(dataset_collection, part_number, **kwargs)
        for phase, s in self._samplers.items():
            self._dataset_indices[phase] = dict(enumerate(s.random_split_indices(part_number)))


class SamplerFactory:
    def __init__(self) -> None:
        self._registry: dict[str, Callable] = {}

    def register(self, name: str, cls: Callable) -> None:
        self._registry[name.lower()] = cls

    def get(self, name: str) -> Callable:
        return self._registry[name.lower()]

    def has(self, name: str) -> bool:
        return name.lower() in self._registry


global_sampler_factory = SamplerFactory()
global_sampler_factory.register("iid", IIDSampler)
global_sampler_factory.register("random", RandomSampler)


def get_dataset_collection_sampler(name: str, dataset_collection: DatasetCollection, part_number: int,
                                   **kwargs: Any) -> DatasetCollectionSampler:
    return global_sampler_factory.get(name)(dataset_collection=dataset_collection, part_number=part_number, **kwargs)
