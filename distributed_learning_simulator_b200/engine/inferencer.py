"""Inferencer: full-phase evaluation (server-side test metric).

Own version of ``cyy_torch_toolbox.inferencer.Inferencer`` with the surface the
reference's server uses (/root/reference/simulation_lib/server/server.py:32-55).
The log line contains ``test in`` + ``accuracy`` + ``NN.NN%`` because the offline
scraper keys on it (simulation_lib/analysis/analyze_log.py:46-52).
"""
import time
from typing import Any

from torch import nn

from .dataset import DatasetCollection
from .executor import HookConfig, ModelExecutor
from .hyper_parameter import HyperParameter
from .log import get_logger
from .ml_type import MachineLearningPhase


class Inferencer(ModelExecutor):
    def __init__(self, model: nn.Module, dataset_collection: DatasetCollection, phase: MachineLearningPhase,
                 hyper_parameter: HyperParameter, hook_config: HookConfig | None = None) -> None:
        super().__init__(model, dataset_collection, phase, hyper_parameter, hook_config)
        self.evaluator: Any = None  # optional hand-written batched evaluator (ops.eval)
        self._reducer: Any = None

    def shard(self, rank: int, world: int, reducer: Any) -> None:
        """Keep every ``world``-th sample of the phase; ``inference`` then sums over ranks."""
        util = self._dc.get_dataset_util(self._phase)
        self._dc.set_subset(self._phase, util.indices[rank::world])
        self._reducer = reducer

    def inference(self, **kwargs: Any) -> None:
        start = time.monotonic()
        metric = self._evaluate(self._phase, reducer=self._reducer)
        self._metric.set_epoch_metrics(1, metric["loss"], metric["accuracy"], float(self.dataset_size))
        if self.hook_config.log_performance_metric:
            get_logger().info("%s%s in %.3f s, %s loss: %e, accuracy: %.2f%%", self._visualizer_prefix,
                              self._phase.value, time.monotonic() - start, self._phase.value, metric["loss"],
                              metric["accuracy"] * 100)
