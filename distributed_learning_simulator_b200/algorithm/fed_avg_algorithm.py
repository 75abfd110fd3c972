"""Weighted federated averaging.

Semantics of /root/reference/simulation_lib/algorithm/fed_avg_algorithm.py:11-110:
dataset-size weights (overridable ``_get_weight``), per-name total weights and the
overridable ``_apply_total_weight``, NaN check, ``end_training`` taken from the first
contributing worker, ``other_data`` merged with a conflict check, a non-accumulating
path (``accumulate=False``) that keeps every worker's model for the Shapley algorithms.

Execution differs: the reference adds each upload into fp64 CPU accumulators with a
Python loop per tensor (fed_avg_algorithm.py:39-52).  Here uploads are ``[P]`` rows in
HBM; when every contribution is such a row the whole round is ONE launch of the
hand-written ``fedavg_reduce`` kernel (ops/reduce.py: fp64 accumulation in registers,
``old + sum(w_i * delta_i) / sum(w_i)``, decode of packed uploads at load time, NaN flag
in the epilogue).  Anything else (custom weights, CPU tensors, partial dicts) takes the
streaming torch path below, which is also the oracle for the kernel tests.
"""
from typing import Any

import torch

from ..engine.log import get_logger
from ..engine.tensor import TensorDict
from ..message import Message, ParameterMessage
from .aggregation_algorithm import AggregationAlgorithm


class _StreamingSum:
    """fp64 running sums ``sum_w weight_w * value_w`` and ``sum_w weight_w`` per tensor name -- the fallback for
    uploads the fused kernel does not take (CPU tensors, custom weights, partial dicts).  A worker's tensors can
    be dropped as soon as they have been added."""

    def __init__(self) -> None:
        self.weighted: TensorDict = {}
        self.weights: dict[str, Any] = {}
        self.dtypes: dict[str, Any] = {}

    def __bool__(self) -> bool:
        return bool(self.weighted)

    def add(self, name: str, value: torch.Tensor, weight: Any) -> None:
        self.dtypes[name] = value.dtype
        term = value.to(dtype=torch.float64) * weight
        if name in self.weighted:
            self.weighted[name] += term
            self.weights[name] = self.weights[name] + weight
        else:
            self.weighted[name] = term
            self.weights[name] = weight

    def finish(self, normalise: Any) -> TensorDict:
        """``normalise(name, weighted_sum, total_weight)`` per tensor, cast back to the upload dtype."""
        result = {name: normalise(name, total, self.weights[name]).to(dtype=self.dtypes[name])
                  for name, total in self.weighted.items()}
        self.__init__()
        if bool(torch.stack([t.isnan().any() for t in result.values()]).any()):
            raise AssertionError("NaN in aggregated parameters")
        return result


class FedAVGAlgorithm(AggregationAlgorithm):
    # subclasses that change _get_weight/_apply_total_weight set this to the name of the
    # fused-kernel variant implementing the same maths, or None for "no fused path"
    fused_variant: str | None = "fedavg"

    def __init__(self) -> None:
        super().__init__()
        self.accumulate: bool = True
        self.use_fused: bool = True
        self.__streamed = _StreamingSum()
        self.__pending: dict[int, Message] = {}   # raw uploads kept for the fused kernel

    # ---- upload handling -------------------------------------------------------------
    def process_worker_data(self, worker_id: int, worker_data: Message | None,
                            old_parameter_dict: TensorDict | None, save_dir: str) -> None:
        if worker_data is None:
            self._skipped_workers.add(worker_id)
            return
        from ..ops import reduce as reduce_ops

        self._old_parameter_dict = old_parameter_dict
        fusable = self.accumulate and self.use_fused and self.fused_variant is not None
        # the two accumulators are never live together: once something has been streamed this round, every later
        # upload is streamed as well (and deferred rows are folded in the moment a non-fusable upload arrives)
        if fusable and not self.__streamed and reduce_ops.can_fuse(worker_data, old_parameter_dict):
            self.__pending[worker_id] = worker_data   # the kernel consumes the raw (delta / packed) row
            self._all_worker_data[worker_id] = worker_data
            return
        self.__stream_pending()
        self.__stream(worker_id, worker_data)

    def __stream(self, worker_id: int, raw: Message) -> None:
        from ..ops import reduce as reduce_ops

        upload = self._process_worker_data(worker_data=reduce_ops.materialize(raw),
                                           old_parameter_dict=self._old_parameter_dict)
        self._all_worker_data[worker_id] = upload
        if self.accumulate and isinstance(upload, ParameterMessage):
            for name, value in upload.parameter.items():
                self.__streamed.add(name, value, self._get_weight(dataset_size=upload.dataset_size, name=name,
                                                                  parameter=value))
            upload.parameter = {}   # the sums hold everything that is needed: free the worker's copy

    def __stream_pending(self) -> None:
        pending, self.__pending = self.__pending, {}
        for worker_id, raw in pending.items():
            self.__stream(worker_id, raw)

    def _get_weight(self, dataset_size: int, name: str, parameter: Any) -> Any:
        assert dataset_size != 0
        return dataset_size

    def _apply_total_weight(self, name: str, parameter: torch.Tensor, total_weight: Any) -> torch.Tensor:
        return parameter / total_weight

    # ---- aggregation -------------------------------------------------------------------
    def aggregate_worker_data(self) -> Message:
        if not self.accumulate:
            return self._aggregate_worker_data(self._all_worker_data)
        if self.__pending:
            from ..ops import reduce as reduce_ops

            pending, self.__pending = self.__pending, {}
            parameter = reduce_ops.fused_aggregate(pending, self._old_parameter_dict, variant=self.fused_variant)
        else:
            assert self.__streamed, "no upload to aggregate"
            parameter = self.__streamed.finish(
                lambda name, total, weight: self._apply_total_weight(name=name, parameter=total, total_weight=weight))
        return ParameterMessage(
            parameter=parameter,
            end_training=next(iter(self._all_worker_data.values())).end_training,
            other_data=self._check_and_reduce_other_data(self._all_worker_data),
        )

    def clear_worker_data(self) -> None:
        super().clear_worker_data()
        self.__pending = {}
        self.__streamed = _StreamingSum()

    @classmethod
    def _aggregate_worker_data(cls, all_worker_data: dict[int, ParameterMessage]) -> ParameterMessage:
        assert all_worker_data
        assert isinstance(next(iter(all_worker_data.values())), ParameterMessage)
        parameter = AggregationAlgorithm.weighted_avg(all_worker_data, AggregationAlgorithm.get_ratios(all_worker_data))
        assert parameter
        return ParameterMessage(parameter=parameter,
                                end_training=next(iter(all_worker_data.values())).end_training,
                                other_data=cls._check_and_reduce_other_data(all_worker_data))

    @classmethod
    def _check_and_reduce_other_data(cls, all_worker_data: dict) -> dict:
        result: dict = {}
        for worker_data in all_worker_data.values():
            for k, v in worker_data.other_data.items():
                if k not in result:
                    result[k] = v
                    continue
                if v != result[k]:
                    get_logger().error("different values on key %s", k)
                    raise RuntimeError(f"different values on key {k}")
        return result
