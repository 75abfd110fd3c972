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
        """The full model ``parameter + delta`` as a ParameterMessage that keeps every other field of this one."""
        base, delta = as_flat(parameter), as_flat(self.delta_parameter)
        if base is not None and delta is not None and base.layout == delta.layout:
            # one fused add over the whole row
            summed: TensorDict = FlatParameter(base.flat + delta.flat.to(base.flat.device), base.layout)
        else:
            summed = {name: (old + self.delta_parameter[name].to(old.device) if name in self.delta_parameter else old)
                      for name, old in parameter.items()}
        carried = {f.name: getattr(self, f.name) for f in fields(self) if f.name != "delta_parameter"}
        return ParameterMessage(parameter=summed, **carried)


def get_message_size(msg: Any) -> int:
    """Payload bytes of a message: every distinct tensor object reachable from it counted once."""
    sizes: dict[int, int] = {}
    for_each_tensor(msg, lambda t: sizes.setdefault(id(t), t.element_size() * t.numel()))
    total = sum(sizes.values())
    if total <= 0:
        raise ValueError("the message carries no tensor")
    return total
