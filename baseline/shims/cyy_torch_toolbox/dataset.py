"""Synthetic vision datasets (same formulas and sizes as the framework under test, so both
arms of the benchmark train on data of identical shape), cached on the CPU like the
reference's ``cache_transforms: cpu``; label-stratified IID sampler."""
import torch

from .ml_type import MachineLearningPhase

SPECS = {
    "MNIST": dict(shape=(1, 28, 28), classes=10, train=60000, test=10000),
    "CIFAR10": dict(shape=(3, 32, 32), classes=10, train=50000, test=10000),
    "CIFAR100": dict(shape=(3, 32, 32), classes=100, train=50000, test=10000),
}
_cache: dict = {}


def _generate(name: str, kwargs: dict) -> dict:
    key = (name, tuple(sorted((k, str(v)) for k, v in kwargs.items())))
    if key in _cache:
        return _cache[key]
    spec = dict(SPECS[name])
    spec["train"] = kwargs.get("train_size", spec["train"])
    spec["test"] = kwargs.get("test_size", spec["test"])
    g = torch.Generator().manual_seed(int(kwargs.get("seed", 20240229)) + sum(map(ord, name)))
    c, h, w = spec["shape"]
    lo = torch.randn(spec["classes"], c, max(h // 4, 1), max(w // 4, 1), generator=g)
    proto = torch.nn.functional.interpolate(lo, size=(h, w), mode="bilinear", align_corners=False)
    proto = proto / proto.flatten(1).std(dim=1).view(-1, 1, 1, 1)
    out = {"spec": spec}
    for phase, n in (("train", spec["train"]), ("test", spec["test"])):
        y = torch.randint(0, spec["classes"], (n,), generator=g)
        x = torch.randn(n, c, h, w, generator=g)
        x.add_(proto[y], alpha=0.6)
        out[phase] = (x, y)
    _cache[key] = out
    return out


class _Phase:
    def __init__(self, x, y, indices) -> None:
        self.x, self.y, self.indices = x, y, indices

    def __len__(self) -> int:
        return int(self.indices.numel())


class DatasetCollection:
    def __init__(self, name: str, dataset_kwargs: dict | None = None) -> None:
        self.name = name
        self.dataset_kwargs = dict(dataset_kwargs or {})
        base = _generate(name, self.dataset_kwargs)
        self.spec = base["spec"]
        x, y = base["train"]
        tx, ty = base["test"]
        half = len(ty) // 2
        self._phases = {
            MachineLearningPhase.Training: _Phase(x, y, torch.arange(len(y))),
            MachineLearningPhase.Validation: _Phase(tx, ty, torch.arange(0, half)),
            MachineLearningPhase.Test: _Phase(tx, ty, torch.arange(half, len(ty))),
        }

    def has_dataset(self, phase) -> bool:
        return phase in self._phases

    def remove_dataset(self, phase) -> None:
        self._phases.pop(phase, None)

    def get_phase(self, phase) -> _Phase:
        return self._phases[phase]

    def get_dataset_size(self, phase) -> int:
        return len(self._phases[phase])

    def get_labels(self) -> set:
        return set(range(self.spec["classes"]))

    def is_mutilabel(self) -> bool:
        return False

    def set_subset(self, phase, indices) -> None:
        p = self._phases[phase]
        self._phases[phase] = _Phase(p.x, p.y, indices)


class ClassificationDatasetCollection(DatasetCollection):
    pass


class _PhaseSampler:
    def __init__(self, phase: _Phase, seed: int) -> None:
        self._phase = phase
        self._g = torch.Generator().manual_seed(seed)

    def split_indices(self, parts: list) -> list:
        labels = self._phase.y[self._phase.indices]
        result = [[] for _ in parts]
        for label in sorted(set(labels.tolist())):
            idx = self._phase.indices[labels == label]
            owners = [i for i, p in enumerate(parts) if p.get(label, 0) > 0]
            if not owners:
                continue
            idx = idx[torch.randperm(idx.numel(), generator=self._g)]
            for chunk, i in zip(torch.tensor_split(idx, len(owners)), owners):
                result[i].append(chunk)
        return [torch.sort(torch.cat(r))[0] if r else torch.empty(0, dtype=torch.long) for r in result]


class DatasetCollectionSampler:
    def __init__(self, dataset_collection, part_number: int, **kwargs) -> None:
        self._dc = dataset_collection
        self._part_number = part_number
        self._dataset_indices: dict = {}
        self._samplers = {phase: _PhaseSampler(dataset_collection.get_phase(phase), i)
                          for i, phase in enumerate(MachineLearningPhase) if dataset_collection.has_dataset(phase)}

    def set_dataset_collection(self, dataset_collection) -> None:
        self._dc = dataset_collection

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_dc"] = None       # the (large) tensors are regenerated from the seed in each process
        state["_samplers"] = {}
        return state

    def sample(self, part_id: int) -> None:
        for phase, parts in self._dataset_indices.items():
            if self._dc.has_dataset(phase):
                self._dc.set_subset(phase, parts[part_id])


class IIDSampler(DatasetCollectionSampler):
    def __init__(self, dataset_collection, part_number: int, **kwargs) -> None:
        super().__init__(dataset_collection, part_number)
        labels = dataset_collection.get_labels()
        for phase, s in self._samplers.items():
            self._dataset_indices[phase] = dict(enumerate(
                s.split_indices([{label: 1 for label in labels} for _ in range(part_number)])))


class _Factory:
    def __init__(self) -> None:
        self._reg = {"iid": IIDSampler}

    def register(self, name: str, cls) -> None:
        self._reg[name.lower()] = cls

    def get(self, name: str):
        return self._reg[name.lower()]


global_sampler_factory = _Factory()


def get_dataset_collection_sampler(name, dataset_collection, part_number, **kwargs):
    return global_sampler_factory.get(name)(dataset_collection=dataset_collection, part_number=part_number, **kwargs)
