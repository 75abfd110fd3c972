"""FedAvg that also computes per-round Shapley values of the workers.

Behaviour of /root/reference/simulation_lib/method/shapley_value/shapley_value_algorithm.py:
11-92: ``accumulate=False`` (every worker's full model is kept), metric ``accuracy``, the
engine is created lazily at round 1 with ``last_round_metric = stat[0].test_accuracy``,
``shapley_values[round]`` / ``shapley_values_S[round]`` snapshots, optional
``choose_best_subset`` aggregation, ``shapley_values.json`` (+``_S``) written on exit.

The utility of a coalition is the test accuracy of the dataset-size-weighted average of
its members' models (:67-76).  The reference builds that average on the CPU and runs one
full test inference per coalition, strictly sequentially.  Here the N uploaded models are
``[P]`` rows in HBM; a coalition model is one launch of the fused reduce kernel, and when
the tester model is supported by ``ops.subset_eval`` many coalitions are evaluated in one
batched launch (the engines call ``prefetch``).
"""
import copy
import json
import os
from typing import Any, Iterable, Type

from ...algorithm.shapley_value import GTGShapleyValue, HierarchicalShapleyValue, MultiRoundShapleyValue
from ...engine.log import get_logger
from ..common_import import AggregationServer, FedAVGAlgorithm, Message
from ...utils import tracing


class ShapleyValueAlgorithm(FedAVGAlgorithm):
    def __init__(self, sv_algorithm_cls: Type, server: AggregationServer, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._server: AggregationServer = server
        self.accumulate = False
        self.metric_type: str = "accuracy"
        self.sv_algorithm: Any = None
        self.sv_algorithm_cls = sv_algorithm_cls
        self.shapley_values: dict = {}
        self.shapley_values_S: dict = {}
        self.subset_evaluations = 0

    @property
    def config(self) -> Any:
        return self._server.config

    @property
    def choose_best_subset(self) -> bool:
        return self.config.algorithm_kwargs.get("choose_best_subset", False)

    def _get_players(self) -> Iterable:
        return sorted(self._all_worker_data.keys())

    def _engine_kwargs(self) -> dict:
        kwargs = {"seed": self.config.reproducible_env_config.seed}
        if "speculative_prefetch" in self.config.algorithm_kwargs:
            kwargs["speculative_prefetch"] = bool(self.config.algorithm_kwargs["speculative_prefetch"])
        return kwargs

    def aggregate_worker_data(self) -> Message:
        if self.sv_algorithm is None:
            assert self._server.round_number == 1
            self.sv_algorithm = self.sv_algorithm_cls(
                players=self._get_players(),
                last_round_metric=self._server.performance_stat[self._server.round_number - 1][
                    f"test_{self.metric_type}"],
                **self._engine_kwargs(),
            )
        assert self.sv_algorithm is not None
        self.sv_algorithm.set_metric_function(self._get_subset_metric)
        self.sv_algorithm.set_batch_metric_function(self._batch_evaluator())
        with tracing.phase("shapley"):
            self.sv_algorithm.compute(round_number=self._server.round_number)
        self.subset_evaluations += self.sv_algorithm.evaluations
        self.sv_algorithm.evaluations = 0
        self.shapley_values[self._server.round_number] = copy.deepcopy(
            self._convert_shapley_values(self.sv_algorithm.shapley_values))
        self.shapley_values_S[self._server.round_number] = self._convert_shapley_values(
            self.sv_algorithm.shapley_values_S)
        if self.choose_best_subset:
            best_subset: set = set(self.shapley_values_S[self._server.round_number].keys())
            if best_subset:
                get_logger().warning("use subset %s", best_subset)
                self._all_worker_data = {k: v for k, v in self._all_worker_data.items() if k in best_subset}
        return super().aggregate_worker_data()

    def _convert_shapley_values(self, shapley_values: dict) -> dict:
        return dict(shapley_values)

    def _get_subset_metric(self, subset: Iterable) -> float:
        assert subset
        members = set(subset)
        worker_data = FedAVGAlgorithm._aggregate_worker_data(
            {k: v for k, v in self._all_worker_data.items() if k in members})
        assert worker_data
        return self._server.get_metric(worker_data, keep_performance_logger=False)[self.metric_type]

    def _batch_evaluator(self) -> Any:
        """A callable evaluating a list of coalitions in one launch, or None."""
        from ...ops import subset_eval

        return subset_eval.make_batch_evaluator(self._server, self._all_worker_data, self.metric_type)

    def exit(self) -> None:
        if self.sv_algorithm is None:
            return
        os.makedirs(self.config.save_dir, exist_ok=True)
        with open(os.path.join(self.config.save_dir, "shapley_values.json"), "wt", encoding="utf8") as f:
            json.dump(self.shapley_values, f)
        if self.choose_best_subset:
            with open(os.path.join(self.config.save_dir, "shapley_values_S.json"), "wt", encoding="utf8") as f:
                json.dump(self.shapley_values_S, f)


class GTGShapleyValueAlgorithm(ShapleyValueAlgorithm):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(GTGShapleyValue, *args, **kwargs)


class MultiRoundShapleyValueAlgorithm(ShapleyValueAlgorithm):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(MultiRoundShapleyValue, *args, **kwargs)


class HierarchicalShapleyValueAlgorithm(ShapleyValueAlgorithm):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(HierarchicalShapleyValue, *args, **kwargs)

    def _engine_kwargs(self) -> dict:
        kw = self.config.algorithm_kwargs
        return {"part_number": kw.get("part_number", 2), "vp_size": kw.get("vp_size")}
