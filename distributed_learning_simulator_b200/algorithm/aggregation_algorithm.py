"""Base of the server-side aggregation algorithms.

API parity with /root/reference/simulation_lib/algorithm/aggregation_algorithm.py:9-96
(``get_ratios``, ``weighted_avg``, ``process_worker_data``, ``aggregate_worker_data``,
``clear_worker_data``, ``exit``; skipped workers recorded on ``None``).  Payloads stay on
the device they arrived on (the reference moves everything to the CPU,
aggregation_algorithm.py:60-67).
"""
from typing import Any

import torch

from ..engine.tensor import TensorDict
from ..message import DeltaParameterMessage, Message, ParameterMessage


class AggregationAlgorithm:
    def __init__(self) -> None:
        self._all_worker_data: dict[int, Message] = {}
        self._skipped_workers: set[int] = set()
        self._old_parameter_dict: TensorDict | None = None

    @classmethod
    def get_ratios(cls, data_dict: dict[int, ParameterMessage], key_name: str | None = None) -> dict[int, float]:
        if key_name is None:
            total_scalar = sum(v.dataset_size for v in data_dict.values())
            return {k: float(v.dataset_size) / float(total_scalar) for k, v in data_dict.items()}
        total_scalar = sum(v.other_data[key_name] for v in data_dict.values())
        return {k: float(v.other_data[key_name]) / float(total_scalar) for k, v in data_dict.items()}

    @classmethod
    def weighted_avg(cls, data_dict: dict[int, ParameterMessage], weight_dict: dict[int, float]) -> TensorDict:
        assert data_dict
        from ..ops import reduce as reduce_ops

        fused = reduce_ops.try_weighted_avg(data_dict, weight_dict)
        if fused is not None:
            return fused
        avg_data: TensorDict = {}
        for worker_id, v in data_dict.items():
            ratio = weight_dict[worker_id]
            assert 0 <= ratio <= 1
            if not avg_data:
                avg_data = {k2: v2 * ratio for (k2, v2) in v.parameter.items()}
            else:
                for k in avg_data:
                    avg_data[k].add_(v.parameter[k].to(avg_data[k].device), alpha=ratio)
        nan = torch.stack([p.isnan().any() for p in avg_data.values()]).any()
        assert not bool(nan), "NaN in aggregated parameters"
        return avg_data

    def _process_worker_data(self, worker_data: Message, old_parameter_dict: TensorDict | None) -> Message:
        match worker_data:
            case DeltaParameterMessage():
                assert old_parameter_dict is not None
                return worker_data.restore(old_parameter_dict)
            case ParameterMessage():
                if old_parameter_dict is not None:
                    worker_data.complete(old_parameter_dict)
                return worker_data
            case Message():
                return worker_data
        raise NotImplementedError(worker_data)

    def process_worker_data(self, worker_id: int, worker_data: Message | None,
                            old_parameter_dict: TensorDict | None, save_dir: str) -> None:
        if worker_data is None:
            self._skipped_workers.add(worker_id)
            return
        self._old_parameter_dict = old_parameter_dict
        self._all_worker_data[worker_id] = self._process_worker_data(
            worker_data=worker_data, old_parameter_dict=old_parameter_dict)

    def aggregate_worker_data(self) -> Any:
        raise NotImplementedError()

    def clear_worker_data(self) -> None:
        self._all_worker_data.clear()
        self._skipped_workers.clear()

    def exit(self) -> None:
        pass
