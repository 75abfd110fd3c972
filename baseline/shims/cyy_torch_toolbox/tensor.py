import dataclasses

import torch


def recursive_tensor_op(data, fun, **kwargs):
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


def tensor_to(data, non_blocking=True, **kwargs):
    return recursive_tensor_op(data, lambda t, **kw: t.to(non_blocking=non_blocking, **kw), **kwargs)


def cat_tensors_to_vector(tensors):
    return torch.cat([t.reshape(-1) for t in tensors])


def decompose_tensor_to_list(shapes, tensor):
    out, bias = [], 0
    for shape in shapes:
        n = 1
        for s in shape:
            n *= s
        out.append(tensor[bias: bias + n].view(*shape))
        bias += n
    return out
