"""Best-model-by-validation hook.

Own version of ``cyy_torch_toolbox.hook.keep_model.KeepModelHook`` as used by the
reference's aggregation worker (/root/reference/simulation_lib/worker/
aggregation_worker.py:62-80,86): after each epoch's validation pass the parameters are
snapshotted if validation accuracy improved; ``best_model["parameter"]`` is what gets
uploaded under IID sampling.  The snapshot is a single flat device copy (one D2D of
``[P]``) when the model lives in the flat arena.
"""
from typing import Any

from .hook import Hook
from .model_util import FlatParameter
from .tensor import tensor_clone


class KeepModelHook(Hook):
    def __init__(self) -> None:
        super().__init__()
        self.keep_best_model: bool = False
        self.__best: dict | None = None

    @property
    def best_model(self) -> dict:
        assert self.__best is not None, "no validation has run yet"
        return self.__best

    def has_best_model(self) -> bool:
        return self.__best is not None

    def clear(self) -> None:
        self.__best = None

    def _after_validation(self, executor: Any, epoch: int, validation_metric: dict, **kwargs: Any) -> None:
        if not self.keep_best_model:
            return
        acc = validation_metric["accuracy"]
        if self.__best is not None and acc <= self.__best["performance_metric"]["accuracy"]:
            return
        util = executor.model_util
        flat = util.flat_parameter
        if flat is not None:
            snap = flat.detach().clone()
            parameter = FlatParameter(snap, util.layout)
        else:
            parameter = tensor_clone(util.get_parameter_dict())
        self.__best = {"epoch": epoch, "parameter": parameter, "performance_metric": dict(validation_metric)}
