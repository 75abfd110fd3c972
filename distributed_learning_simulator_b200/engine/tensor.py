"""Tensor helpers over nested containers.

Own implementations of the ``cyy_torch_toolbox.tensor`` helpers the reference consumes
(tensor_to, cat_tensors_to_vector, decompose_tensor_to_list, recursive_tensor_op; call
sites: /root/reference/simulation_lib/algorithm/aggregation_algorithm.py:3,
util/model_cache.py:5, worker/gradient_worker.py:6-7, message.py:6).
"""
import dataclasses
from typing import Any, Callable, Iterable, Sequence

import torch

TensorDict = dict[str, torch.Tensor]


def recursive_tensor_op(data: Any, fun: Callable, **kwargs: Any) -> Any:
    """Apply ``fun`` to every tensor inside lists/tuples/dicts/sets/dataclasses."""
    if isinstance(data, torch.Tensor):
        return fun(data, **kwargs)
    if isinstance(data, dict):
        return {k: recursive_tensor_op(v, fun, **kwargs) for k, v in data.items()}
    if isinstance(data, list):
        return [recursive_tensor_op(v, fun, **kwargs) for v in data]
    if isinstance(data, tuple):
        return tuple(recursive_tensor_op(v, fun, **kwargs) for v in data)
    if dataclasses.is_dataclass(data) and not isinstance(data, type):
        for f in dataclasses.fields(data):
            setattr(data, f.name, recursive_tensor_op(getattr(data, f.name), fun, **kwargs))
        return data
    return data


def for_each_tensor(data: Any, fun: Callable) -> None:
    """Visit every tensor without rebuilding (or replacing) any container."""
    if isinstance(data, torch.Tensor):
        fun(data)
    elif isinstance(data, dict):
        flat = getattr(data, "flat", None)
        if isinstance(flat, torch.Tensor):
            fun(flat)  # flat payloads: the views share this storage
        else:
            for v in data.values():
                for_each_tensor(v, fun)
        extra = getattr(data, "partial_flat", None)
        if isinstance(extra, torch.Tensor):
            fun(extra)
    elif isinstance(data, (list, tuple, set)):
        for v in data:
            for_each_tensor(v, fun)
    elif dataclasses.is_dataclass(data) and not isinstance(data, type):
        for f in dataclasses.fields(data):
            for_each_tensor(getattr(data, f.name), fun)


def tensor_to(data: Any, non_blocking: bool = True, **kwargs: Any) -> Any:
    def fun(t: torch.Tensor, **kw: Any) -> torch.Tensor:
        return t.to(non_blocking=non_blocking, **kw)

    return recursive_tensor_op(data, fun, **kwargs)


def tensor_clone(data: Any, detach: bool = True) -> Any:
    def fun(t: torch.Tensor) -> torch.Tensor:
        return t.detach().clone() if detach else t.clone()

    return recursive_tensor_op(data, fun)


def cat_tensors_to_vector(tensors: Iterable[torch.Tensor]) -> torch.Tensor:
    return torch.cat([t.reshape(-1) for t in tensors])


def decompose_tensor_to_list(shapes: Sequence[torch.Size], tensor: torch.Tensor) -> list[torch.Tensor]:
    result = []
    bias = 0
    for shape in shapes:
        n = 1
        for s in shape:
            n *= s
        result.append(tensor[bias : bias + n].view(*shape))
        bias += n
    assert bias == tensor.numel()
    return result


def get_tensor_bytes(data: Any) -> int:
    cnt = 0

    def count(t: torch.Tensor) -> torch.Tensor:
        nonlocal cnt
        cnt += t.element_size() * t.numel()
        return t

    recursive_tensor_op(data, count)
    return cnt
