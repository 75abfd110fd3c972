"""Wire messages between workers and the server.

Same dataclass family and field semantics as
/root/reference/simulation_lib/message.py:10-62 (``Message``, ``ParameterMessageBase``,
``ParameterMessage.complete``, ``ParameterFileMessage``, ``DeltaParameterMessage.restore``,
``get_message_size``).  Differences, by design:

* ``restore`` does not deep-copy the old model (the reference does, message.py:42);
  it builds ``old + delta`` out of place, one fused add per tensor.
* A message may carry ``flat`` payloads (``FlatParameter``): a ``[P]`` row of a device
  arena plus its layout.  The dict view is materialised lazily as zero-copy tensor views,
  so code written against the ``TensorDict`` API keeps working while the aggregation
  kernels consume the flat row directly.
"""
from dataclasses import dataclass, field, fields
from typing import Any

import torch

from .engine.model_util import FlatParameter, ParameterLayout, as_flat  # noqa: F401
from .engine.tensor import TensorDict, for_each_tensor


@dataclass(kw_only=True)
class Message:
    other_data: dict = field(default_factory=lambda: {})
    in_round: bool = False
    end_training: bool = False


@dataclass(kw_only=True)
class ParameterMessageBase(Message):
    dataset_size: int = 0


@dataclass(kw_only=True)
class ParameterMessage(ParameterMessageBase):
    parameter: TensorDict

    def complete(self, other_parameter: TensorDict) -> None:
        for k, v in other_parameter.items():
            if k not in self.parameter:
                self.parameter[k] = v


@dataclass(kw_only=True)
class ParameterFileMessage(ParameterMessageBase):
    path: str


@dataclass(kw_only=True)
class DeltaParameterMessage(ParameterMessageBase):
    delta_parameter: TensorDict

    def restore(self, parameter: TensorDict) -> ParameterMessage:
        old_flat, delta_flat = as_flat(parameter), as_flat(self.delta_parameter)
        if old_flat is not None and delta_flat is not None and old_flat.layout == delta_flat.layout:
            new_parameter: TensorDict = FlatParameter(
                old_flat.flat + delta_flat.flat.to(old_flat.flat.device), old_flat.layout)
        else:
            new_parameter = {}
        for k, v in ([] if new_parameter else parameter.items()):
            d = self.delta_parameter.get(k)
            new_parameter[k] = v + d.to(v.device) if d is not None else v
        msg = ParameterMessage(parameter=new_parameter)
        for f in fields(self):
            if f.name != "delta_parameter":
                setattr(msg, f.name, getattr(self, f.name))
        msg.parameter = new_parameter
        return msg


def get_message_size(msg: Any) -> int:
    cnt: int = 0
    seen: set[int] = set()

    def count(data: torch.Tensor) -> None:
        nonlocal cnt
        if id(data) not in seen:
            seen.add(id(data))
            cnt += data.element_size() * data.numel()

    for_each_tensor(msg, count)
    assert cnt > 0
    return cnt
