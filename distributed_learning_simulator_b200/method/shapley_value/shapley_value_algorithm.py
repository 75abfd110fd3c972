"""FedAvg that also attributes each round's gain to the workers (Shapley values).

What /root/reference/simulation_lib/method/shapley_value/shapley_value_algorithm.py:11-92 provides is kept:
every worker's full model is retained (``accumulate=False``), the utility of a coalition is the test
``accuracy`` of the dataset-size-weighted average of its members, the engine is created at round 1 from the
round-0 test metric, ``shapley_values[round]`` / ``shapley_values_S[round]`` are recorded, ``choose_best_subset``
restricts the round's aggregate to the best coalition, and ``exit`` writes ``shapley_values.json`` (+ ``_S``).

How a utility is computed is this framework's own: the uploads are ``[P]`` rows in HBM, a coalition model is one
launch of the fused reduce kernel, and when ``ops.subset_eval`` supports the tester model many coalitions are
evaluated per launch -- the engines ask for whole batches through ``set_batch_metric_function``.
"""
import json
import os
from typing import Any, Iterable, Type

from ...algorithm.shapley_value import GTGShapleyValue, HierarchicalShapleyValue, MultiRoundShapleyValue
from ...engine.log import get_logger
from ...utils import tracing
from ..common_import import AggregationServer, FedAVGAlgorithm, Message


class _CoalitionUtility:
    """v(S) for the uploads of the current round."""

    def __init__(self, server: AggregationServer, uploads: dict, metric_type: str) -> None:
        self.server, self.uploads, self.metric_type = server, uploads, metric_type

    def __call__(self, subset: Iterable) -> float:
        members = set(subset)
        if not members:
            raise ValueError("the empty coalition has no model")
        model = FedAVGAlgorithm._aggregate_worker_data({w: m for w, m in self.uploads.items() if w in members})
        return self.server.get_metric(model, keep_performance_logger=False)[self.metric_type]

    def batched(self) -> Any:
        """A callable that evaluates a list of coalitions in one launch, or None when the model has no batched path."""
        from ...ops import subset_eval

        return subset_eval.make_batch_evaluator(self.server, self.uploads, self.metric_type)


class ShapleyValueAlgorithm(FedAVGAlgorithm):
    def __init__(self, sv_algorithm_cls: Type, server: AggregationServer, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.accumulate = False          # coalitions need every worker's own model
        self.metric_type = "accuracy"
        self._server = server
        self.sv_algorithm_cls = sv_algorithm_cls
        self.sv_algorithm: Any = None
        self.shapley_values: dict = {}
        self.shapley_values_S: dict = {}
        self.subset_evaluations = 0

    @property
    def config(self) -> Any:
        return self._server.config

    @property
    def choose_best_subset(self) -> bool:
        return bool(self.config.algorithm_kwargs.get("choose_best_subset", False))

    # ---- engine ------------------------------------------------------------------------------------------
    def _get_players(self) -> Iterable:
        return sorted(self._all_worker_data)

    def _engine_kwargs(self) -> dict:
        options = {"seed": self.config.reproducible_env_config.seed}
        if "speculative_prefetch" in self.config.algorithm_kwargs:
            options["speculative_prefetch"] = bool(self.config.algorithm_kwargs["speculative_prefetch"])
        return options

    def __engine(self) -> Any:
        if self.sv_algorithm is None:
            round_number = self._server.round_number
            if round_number != 1:
                raise RuntimeError("the Shapley engine starts at round 1 (it needs the round-0 metric)")
            baseline = self._server.performance_stat[round_number - 1][f"test_{self.metric_type}"]
            self.sv_algorithm = self.sv_algorithm_cls(players=self._get_players(), last_round_metric=baseline,
                                                      **self._engine_kwargs())
        return self.sv_algorithm

    def _convert_shapley_values(self, shapley_values: dict) -> dict:
        return dict(shapley_values)

    # ---- one round ---------------------------------------------------------------------------------------
    def aggregate_worker_data(self) -> Message:
        engine = self.__engine()
        round_number = self._server.round_number
        utility = _CoalitionUtility(self._server, self._all_worker_data, self.metric_type)
        engine.set_metric_function(utility)
        engine.set_batch_metric_function(utility.batched())
        with tracing.phase("shapley"):
            engine.compute(round_number=round_number)
        self.subset_evaluations += engine.evaluations
        engine.evaluations = 0
        self.shapley_values[round_number] = self._convert_shapley_values(engine.shapley_values)
        self.shapley_values_S[round_number] = self._convert_shapley_values(engine.shapley_values_S)
        best = set(self.shapley_values_S[round_number]) if self.choose_best_subset else set()
        if best:
            get_logger().warning("use subset %s", best)
            self._all_worker_data = {w: m for w, m in self._all_worker_data.items() if w in best}
        return super().aggregate_worker_data()

    # kept for subclasses / callers that evaluate a coalition directly
    def _get_subset_metric(self, subset: Iterable) -> float:
        return _CoalitionUtility(self._server, self._all_worker_data, self.metric_type)(subset)

    def exit(self) -> None:
        if self.sv_algorithm is None:
            return
        out_dir = self.config.save_dir
        os.makedirs(out_dir, exist_ok=True)
        outputs = {"shapley_values.json": self.shapley_values}
        if self.choose_best_subset:
            outputs["shapley_values_S.json"] = self.shapley_values_S
        for name, table in outputs.items():
            with open(os.path.join(out_dir, name), "wt", encoding="utf8") as f:
                json.dump(table, f)


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
        options = self.config.algorithm_kwargs
        return {"part_number": options.get("part_number", 2), "vp_size": options.get("vp_size")}
