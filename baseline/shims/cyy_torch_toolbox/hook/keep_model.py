from . import Hook


class KeepModelHook(Hook):
    def __init__(self) -> None:
        super().__init__()
        self.keep_best_model = False
        self._best = None

    @property
    def best_model(self) -> dict:
        assert self._best is not None
        return self._best

    def clear(self) -> None:
        self._best = None

    def _after_validation(self, executor, epoch, validation_metric, **kwargs) -> None:
        if not self.keep_best_model:
            return
        if self._best is not None and validation_metric["accuracy"] <= self._best["performance_metric"]["accuracy"]:
            return
        self._best = {"epoch": epoch, "performance_metric": dict(validation_metric),
                      "parameter": {k: v.clone() for k, v in executor.model_util.get_parameter_dict().items()}}
