"""Base of the server-side aggregation algorithms.

Interface of /root/reference/simulation_lib/algorithm/aggregation_algorithm.py:9-96 (``get_ratios``,
``weighted_avg``, ``process_worker_data``, ``aggregate_worker_data``, ``clear_worker_data``, ``exit``; a ``None``
upload marks a skipped worker).  Two things differ by design: uploads stay on the device they arrived on (the
reference moves every tensor to the CPU first, :60-67), and an upload is *normalised* exactly once, on arrival,
into "a full model" -- deltas are applied to the previous global model, partial models are completed from it --
so that the subclasses only ever average complete models.
"""
from typing import Any, Mapping

import torch

from ..engine.tensor import TensorDict
from ..message import DeltaParameterMessage, Message, ParameterMessage


def _shares(amounts: Mapping[int, float]) -> dict[int, float]:
    total = float(sum(amounts.values()))
    return {worker: float(amount) / total for worker, amount in amounts.items()}


class AggregationAlgorithm:
    def __init__(self) -> None:
        self._all_worker_data: dict[int, Message] = {}
        self._skipped_workers: set[int] = set()
        self._old_parameter_dict: TensorDict | None = None

    # ---- weights -----------------------------------------------------------------------------------------
    @classmethod
    def get_ratios(cls, data_dict: dict[int, ParameterMessage], key_name: str | None = None) -> dict[int, float]:
        """Each worker's share of the total dataset size, or of ``other_data[key_name]`` when a key is given."""
        if key_name is None:
            return _shares({worker: msg.dataset_size for worker, msg in data_dict.items()})
        return _shares({worker: msg.other_data[key_name] for worker, msg in data_dict.items()})

    @classmethod
    def weighted_avg(cls, data_dict: dict[int, ParameterMessage], weight_dict: dict[int, float]) -> TensorDict:
        """``sum_w weight[w] * parameter[w]``: one launch of the fused reduce kernel when the uploads are flat rows
        in HBM, a tensor-by-tensor accumulation otherwise.  NaNs are an error either way."""
        if not data_dict:
            raise ValueError("nothing to average")
        from ..ops import reduce as reduce_ops

        fused = reduce_ops.try_weighted_avg(data_dict, weight_dict)
        if fused is not None:
            return fused
        result: TensorDict = {}
        for worker, msg in data_dict.items():
            weight = weight_dict[worker]
            if not 0 <= weight <= 1:
                raise ValueError(f"weight {weight} of worker {worker} is not a share")
            for name, tensor in msg.parameter.items():
                if name in result:
                    result[name].add_(tensor.to(result[name].device), alpha=weight)
                else:
                    result[name] = tensor * weight
        if bool(torch.stack([t.isnan().any() for t in result.values()]).any()):
            raise AssertionError("NaN in aggregated parameters")
        return result

    # ---- uploads -----------------------------------------------------------------------------------------
    def _process_worker_data(self, worker_data: Message, old_parameter_dict: TensorDict | None) -> Message:
        """Normalises one upload against the previous global model."""
        if isinstance(worker_data, DeltaParameterMessage):
            if old_parameter_dict is None:
                raise ValueError("a delta upload needs the previous global model")
            return worker_data.restore(old_parameter_dict)
        if isinstance(worker_data, ParameterMessage):
            if old_parameter_dict is not None:
                worker_data.complete(old_parameter_dict)
            return worker_data
        if isinstance(worker_data, Message):
            return worker_data
        raise NotImplementedError(f"unknown upload {type(worker_data).__name__}")

    def process_worker_data(self, worker_id: int, worker_data: Message | None,
                            old_parameter_dict: TensorDict | None, save_dir: str) -> None:
        if worker_data is None:
            self._skipped_workers.add(worker_id)
            return
        self._old_parameter_dict = old_parameter_dict
        self._all_worker_data[worker_id] = self._process_worker_data(worker_data, old_parameter_dict)

    def aggregate_worker_data(self) -> Any:
        raise NotImplementedError()

    def clear_worker_data(self) -> None:
        self._all_worker_data.clear()
        self._skipped_workers.clear()

    def exit(self) -> None:
        """Called once when the server shuts down."""
