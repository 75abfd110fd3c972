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


class FedAVGAlgorithm(AggregationAlgorithm):
    # subclasses that change _get_weight/_apply_total_weight set this to the name of the
    # fused-kernel variant implementing the same maths, or None for "no fused path"
    fused_variant: str | None = "fedavg"

    def __init__(self) -> None:
        super().__init__()
        self.accumulate: bool = True
        self.__dtypes: dict[str, Any] = {}
        self.__total_weights: dict[str, Any] = {}
        self.__parameter: TensorDict = {}
        self.__pending: dict[int, Message] = {}   # raw uploads kept for the fused kernel
        self.use_fused: bool = True

    # ---- upload handling -------------------------------------------------------------
    def process_worker_data(self, worker_id: int, worker_data: Message | None,
                            old_parameter_dict: TensorDict | None, save_dir: str) -> None:
        if worker_data is None:
            self._skipped_workers.add(worker_id)
            return
        self._old_parameter_dict = old_parameter_dict
        if self.accumulate and self.use_fused and self.fused_variant is not None:
            from ..ops import reduce as reduce_ops

            # only while nothing went down the streaming path this round: the two accumulators are
            # never live together (a fusable upload after a non-fusable one is streamed as well)
            if not self.__parameter and reduce_ops.can_fuse(worker_data, old_parameter_dict):
                # defer: the kernel consumes the raw (possibly delta / packed) rows
                self.__pending[worker_id] = worker_data
                self._all_worker_data[worker_id] = worker_data
                return
        self.__flush_pending()
        from ..ops import reduce as reduce_ops

        worker_data = reduce_ops.materialize(worker_data)  # arena / packed payloads -> local tensors
        worker_data = self._process_worker_data(worker_data=worker_data, old_parameter_dict=old_parameter_dict)
        self._all_worker_data[worker_id] = worker_data
        if not self.accumulate or not isinstance(worker_data, ParameterMessage):
            return
        self.__accumulate(worker_data)
        # release to reduce memory pressure
        worker_data.parameter = {}

    def __accumulate(self, worker_data: ParameterMessage) -> None:
        for k, v in worker_data.parameter.items():
            self.__dtypes[k] = v.dtype
            weight = self._get_weight(dataset_size=worker_data.dataset_size, name=k, parameter=v)
            tmp = v.to(dtype=torch.float64) * weight
            if k not in self.__parameter:
                self.__parameter[k] = tmp
            else:
                self.__parameter[k] += tmp
            if k not in self.__total_weights:
                self.__total_weights[k] = weight
            else:
                self.__total_weights[k] = self.__total_weights[k] + weight

    def __flush_pending(self) -> None:
        """A non-fusable upload arrived: fold deferred rows into the streaming accumulators."""
        pending, self.__pending = self.__pending, {}
        for worker_id, raw in pending.items():
            from ..ops import reduce as reduce_ops

            raw = reduce_ops.materialize(raw)
            msg = self._process_worker_data(worker_data=raw, old_parameter_dict=self._old_parameter_dict)
            self._all_worker_data[worker_id] = msg
            if isinstance(msg, ParameterMessage):
                self.__accumulate(msg)
                msg.parameter = {}

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
            assert self.__parameter
            parameter = self.__parameter
            self.__parameter = {}
            for k, v in parameter.items():
                parameter[k] = self._apply_total_weight(
                    name=k, parameter=v, total_weight=self.__total_weights[k]).to(dtype=self.__dtypes[k])
            nan = torch.stack([p.isnan().any() for p in parameter.values()]).any()
            assert not bool(nan), "NaN in aggregated parameters"
            self.__total_weights = {}
        return ParameterMessage(
            parameter=parameter,
            end_training=next(iter(self._all_worker_data.values())).end_training,
            other_data=self._check_and_reduce_other_data(self._all_worker_data),
        )

    def clear_worker_data(self) -> None:
        super().clear_worker_data()
        self.__pending = {}
        self.__parameter = {}
        self.__total_weights = {}

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
