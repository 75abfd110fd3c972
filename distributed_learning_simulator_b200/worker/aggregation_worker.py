"""Worker that uploads its model -- or its change since the last global model -- at an aggregation point.

Protocol of /root/reference/simulation_lib/worker/aggregation_worker.py:16-144: aggregation at
``AFTER_EXECUTE`` unless a method moves it, delta uploads (``_send_parameter_diff``), the best model by
validation accuracy under IID sampling, the initial model from the server, ``end_training``, optional
learning-rate reuse, a per-round model cache, and the skip-round handshake of unselected workers (``None`` in,
``None`` back).  Not reproduced: after that handshake the reference falls through to ``None.end_training``
(SURVEY.md defect D1); here the worker simply keeps waiting.

Everything that travels is a flat ``[P]`` row in HBM (``FlatParameter``): a delta is one fused subtraction on
the worker's stream, nothing visits the host.
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
from ..utils import tracing
from .client import Client

_BEST_MODEL_HOOK = "keep_model_hook"


class AggregationWorker(Client):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._aggregation_time: ExecutorHookPoint = ExecutorHookPoint.AFTER_EXECUTE
        self._reuse_learning_rate = False
        self._send_parameter_diff = True
        self._model_cache = ModelCache()
        self.__upload_best_model = False

    # ---- set-up --------------------------------------------------------------------------------------
    def _before_training(self) -> None:
        super()._before_training()
        datasets = self.trainer.dataset_collection
        datasets.remove_dataset(phase=MachineLearningPhase.Test)  # the server owns the test split
        if self.config.dataset_sampling == "iid" and datasets.has_dataset(MachineLearningPhase.Validation):
            self.enable_choose_model_by_validation()
        else:
            datasets.remove_dataset(phase=MachineLearningPhase.Validation)  # nothing would read its metric
        if self.config.distribute_init_parameters:
            self._get_result_from_server()  # the initial global model
            if self._stopped():
                return
        self._register_aggregation()

    def _register_aggregation(self) -> None:
        """(Re-)installs the upload hook at ``_aggregation_time``; methods call this again after moving it."""
        def upload(**hook_kwargs: Any) -> None:
            self._aggregation(sent_data=self._get_sent_data(), **hook_kwargs)

        self.trainer.remove_named_hook(name="aggregation")
        self.trainer.append_named_hook(self._aggregation_time, "aggregation", upload)

    # ---- best model by validation ----------------------------------------------------------------------
    @property
    def best_model_hook(self) -> KeepModelHook | None:
        if not self.trainer.has_hook_obj(_BEST_MODEL_HOOK):
            return None
        hook = self.trainer.get_hook(_BEST_MODEL_HOOK)
        assert isinstance(hook, KeepModelHook)
        return hook

    def enable_choose_model_by_validation(self) -> None:
        self.__upload_best_model = True
        if self.best_model_hook is None:
            assert self.trainer.dataset_collection.has_dataset(phase=MachineLearningPhase.Validation)
            hook = KeepModelHook()
            hook.keep_best_model = True
            self.trainer.append_hook(hook, _BEST_MODEL_HOOK)

    def disable_choose_model_by_validation(self) -> None:
        self.__upload_best_model = False
        if self.best_model_hook is not None:
            self.trainer.remove_hook(_BEST_MODEL_HOOK)

    # ---- upload ----------------------------------------------------------------------------------------
    def _get_parameter_snapshot(self) -> TensorDict:
        """The current parameters, detached from the arena the trainer keeps writing."""
        util = self.trainer.model_util
        if util.flat_parameter is not None:
            return FlatParameter(util.flat_parameter.detach().clone(), util.layout)
        return {name: value.clone() for name, value in util.get_parameter_dict().items()}

    def __model_to_upload(self) -> TensorDict:
        hook = self.best_model_hook if self.__upload_best_model else None
        if hook is not None and hook.has_best_model():
            return hook.best_model["parameter"]
        return self._get_parameter_snapshot()

    def _get_sent_data(self) -> ParameterMessageBase:
        model = self.__model_to_upload()
        size = self.trainer.dataset_size
        if self._send_parameter_diff:
            return DeltaParameterMessage(dataset_size=size, delta_parameter=self._model_cache.get_parameter_diff(model))
        return ParameterMessage(dataset_size=size, parameter=model)

    def _aggregation(self, sent_data: Message, **kwargs: Any) -> None:
        with tracing.phase("upload"):
            self.send_data_to_server(sent_data)
        self._offload_from_device()
        self._get_result_from_server()

    def _offload_from_device(self) -> None:
        if self.config.limited_resource:
            self._model_cache.save()
        hook = self.best_model_hook
        if hook is not None:
            hook.clear()
        super()._offload_from_device()

    # ---- answers ---------------------------------------------------------------------------------------
    def _load_result_from_server(self, result: Message) -> None:
        if result.end_training:
            self._force_stop = True
            raise StopExecutingException()
        if "resume_round" in result.other_data:  # the server continues a check-pointed run
            self._round_num = int(result.other_data["resume_round"])
        path = os.path.join(self.config.save_dir, "aggregated_model", f"round_{self._round_num}.pk")
        if isinstance(result, ParameterMessage):
            self._model_cache.cache_parameter_dict(result.parameter, path=path)
        elif isinstance(result, DeltaParameterMessage):
            self._model_cache.add_parameter_diff(result.delta_parameter, path=path)
        else:
            raise NotImplementedError(f"cannot load a {type(result).__name__}")
        load_parameters(trainer=self.trainer, parameter_dict=self._model_cache.parameter_dict,
                        reuse_learning_rate=self._reuse_learning_rate)

    def _get_result_from_server(self) -> None:
        """Blocks until the server sends a model.  ``None`` means "not selected this round": acknowledge, count the
        round and wait for the next answer."""
        while True:
            answer = super()._get_data_from_server()
            if answer is not None:
                self._load_result_from_server(result=answer)
                return
            get_logger().debug("skip round %s", self._round_num)
            self._round_num += 1
            self.send_data_to_server(None)
            if self._stopped():
                return
