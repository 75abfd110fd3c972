"""Worker that uploads its model (or model delta) once per aggregation point.

Behaviour parity with /root/reference/simulation_lib/worker/aggregation_worker.py:16-144:
aggregation at ``AFTER_EXECUTE`` by default, ``_send_parameter_diff`` (delta uploads),
best-model-by-validation under IID sampling (``KeepModelHook``), initial model from the
server, ``_load_result_from_server`` with ``end_training`` handling and optional
learning-rate reuse, per-round model cache, the skip-round protocol for unselected
workers.  Fixed: after acknowledging a skipped round the worker keeps waiting instead of
falling through to ``None.end_training`` (SURVEY.md defect D1).

Uploads are flat ``[P]`` rows in HBM (``FlatParameter``): the delta is one fused
subtraction on the worker's stream and nothing is copied to the host.
"""
import os
from typing import Any

from ..engine.keep_model import KeepModelHook
from ..engine.log import get_logger
from ..engine.ml_type import ExecutorHookPoint, MachineLearningPhase, StopExecutingException
from ..engine.tensor import TensorDict
from ..message import DeltaParameterMessage, FlatParameter, Message, ParameterMessage, ParameterMessageBase
from ..util.model import load_parameters
from ..util.model_cache import ModelCache
from .client import Client


class AggregationWorker(Client):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._aggregation_time: ExecutorHookPoint = ExecutorHookPoint.AFTER_EXECUTE
        self._reuse_learning_rate: bool = False
        self.__choose_model_by_validation: bool = False
        self._send_parameter_diff: bool = True
        self._model_cache: ModelCache = ModelCache()

    def _before_training(self) -> None:
        super()._before_training()
        self.trainer.dataset_collection.remove_dataset(phase=MachineLearningPhase.Test)
        if self.config.dataset_sampling == "iid" and \
                self.trainer.dataset_collection.has_dataset(MachineLearningPhase.Validation):
            self.enable_choose_model_by_validation()
        if not self.__choose_model_by_validation:
            # Skip Validation to speed up training
            self.trainer.dataset_collection.remove_dataset(phase=MachineLearningPhase.Validation)
        # load initial parameters
        if self.config.distribute_init_parameters:
            self._get_result_from_server()
            if self._stopped():
                return
        self._register_aggregation()

    def _register_aggregation(self) -> None:
        get_logger().debug("use aggregation_time %s", self._aggregation_time)
        self.trainer.remove_named_hook(name="aggregation")

        def __aggregation_impl(**kwargs: Any) -> None:
            self._aggregation(sent_data=self._get_sent_data(), **kwargs)

        self.trainer.append_named_hook(self._aggregation_time, "aggregation", __aggregation_impl)

    def _aggregation(self, sent_data: Message, **kwargs: Any) -> None:
        from ..utils import tracing

        with tracing.phase("upload"):
            self.send_data_to_server(sent_data)
        self._offload_from_device()
        self._get_result_from_server()

    def enable_choose_model_by_validation(self) -> None:
        self.__choose_model_by_validation = True
        if self.trainer.has_hook_obj("keep_model_hook"):
            return
        hook = KeepModelHook()
        hook.keep_best_model = True
        assert self.trainer.dataset_collection.has_dataset(phase=MachineLearningPhase.Validation)
        self.trainer.append_hook(hook, "keep_model_hook")

    def disable_choose_model_by_validation(self) -> None:
        self.__choose_model_by_validation = False
        if self.best_model_hook is not None:
            self.trainer.remove_hook("keep_model_hook")

    @property
    def best_model_hook(self) -> KeepModelHook | None:
        if not self.trainer.has_hook_obj("keep_model_hook"):
            return None
        hook = self.trainer.get_hook("keep_model_hook")
        assert isinstance(hook, KeepModelHook)
        return hook

    def _get_parameter_snapshot(self) -> TensorDict:
        """Current parameters, decoupled from the live arena the trainer keeps updating."""
        util = self.trainer.model_util
        flat = util.flat_parameter
        if flat is not None:
            return FlatParameter(flat.detach().clone(), util.layout)
        return {k: v.clone() for k, v in util.get_parameter_dict().items()}

    def _get_sent_data(self) -> ParameterMessageBase:
        parameter: TensorDict
        if self.__choose_model_by_validation and self.best_model_hook is not None \
                and self.best_model_hook.has_best_model():
            get_logger().debug("use best model")
            parameter = self.best_model_hook.best_model["parameter"]
        else:
            parameter = self._get_parameter_snapshot()
        if self._send_parameter_diff:
            return DeltaParameterMessage(dataset_size=self.trainer.dataset_size,
                                         delta_parameter=self._model_cache.get_parameter_diff(parameter))
        return ParameterMessage(dataset_size=self.trainer.dataset_size, parameter=parameter)

    def _load_result_from_server(self, result: Message) -> None:
        if result.end_training:
            self._force_stop = True
            raise StopExecutingException()
        if "resume_round" in result.other_data:  # the server continues a check-pointed run
            self._round_num = int(result.other_data["resume_round"])
        model_path = os.path.join(self.config.save_dir, "aggregated_model", f"round_{self._round_num}.pk")
        match result:
            case ParameterMessage():
                self._model_cache.cache_parameter_dict(result.parameter, path=model_path)
            case DeltaParameterMessage():
                self._model_cache.add_parameter_diff(result.delta_parameter, path=model_path)
            case _:
                raise NotImplementedError()
        load_parameters(trainer=self.trainer, parameter_dict=self._model_cache.parameter_dict,
                        reuse_learning_rate=self._reuse_learning_rate)

    def _offload_from_device(self) -> None:
        if self.config.limited_resource:
            self._model_cache.save()
        if self.best_model_hook is not None:
            self.best_model_hook.clear()
        super()._offload_from_device()

    def _get_result_from_server(self) -> None:
        while True:
            result = super()._get_data_from_server()
            get_logger().debug("get result from server %s", type(result))
            if result is None:
                get_logger().debug("skip round %s", self._round_num)
                self._round_num += 1
                self.send_data_to_server(None)
                if self._stopped():
                    return
                continue
            self._load_result_from_server(result=result)
            return
