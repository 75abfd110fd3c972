"""FedOBD worker (/root/reference/simulation_lib/method/fed_obd/worker.py:12-74).

Stage one: after every local round, upload only the most-changed blocks (quantized by the
endpoint).  Stage two (entered when a result carries ``phase_two``): learning-rate reuse,
delta uploads of the whole model after every epoch with ``check_acc``, ``end_training`` on
the last epoch.

Fixed (SURVEY.md defect D4): the reference puts *deltas* of the kept blocks into a
``ParameterMessage`` that the server then completes with *absolute* old values, so kept
blocks average to a mean delta.  Here the upload carries the kept blocks' absolute
parameters, dropped blocks fall back to the previous global model, then FedAvg."""
from typing import Any

from ...engine.log import get_logger
from ...engine.ml_type import ExecutorHookPoint
from ..common_import import AggregationWorker, Message, ParameterMessage, QuantClientEndpoint
from .obd_algorithm import OpportunisticBlockDropoutAlgorithm
from .phase import Phase


class FedOBDWorker(AggregationWorker, OpportunisticBlockDropoutAlgorithm):
    def __init__(self, *args: Any, **kwargs: Any) -> None:
        AggregationWorker.__init__(self, *args, **kwargs)
        OpportunisticBlockDropoutAlgorithm.__init__(
            self, dropout_rate=self.config.algorithm_kwargs["dropout_rate"], worker_id=self.worker_id)
        assert isinstance(self._endpoint, QuantClientEndpoint)
        self._endpoint.dequant_server_data = True
        self._send_parameter_diff = False
        self.__phase = Phase.STAGE_ONE
        self.__end_training = False

    @property
    def phase(self) -> Phase:
        return self.__phase

    def _load_result_from_server(self, result: Message) -> None:
        switch = "phase_two" in result.other_data and self.__phase == Phase.STAGE_ONE
        if switch:
            assert isinstance(result, ParameterMessage)
            self.__phase = Phase.STAGE_TWO
            get_logger().warning("switch to phase 2")
            self._reuse_learning_rate = True
            self._send_parameter_diff = True
            self.disable_choose_model_by_validation()
            self.trainer.hyper_parameter.epoch = self.config.algorithm_kwargs["second_phase_epoch"]
            self.config.round = self._round_num + 1
            get_logger().warning("change epoch to %s", self.trainer.hyper_parameter.epoch)
            self._aggregation_time = ExecutorHookPoint.AFTER_EPOCH
            self._register_aggregation()
        super()._load_result_from_server(result=result)
        if switch:
            # stage two is one long train() with its own schedule; inside it the schedule is kept
            self.trainer.remove_optimizer()

    def _aggregation(self, sent_data: Message, **kwargs: Any) -> None:
        if self.__phase == Phase.STAGE_TWO:
            executor = kwargs["executor"]
            if kwargs["epoch"] == executor.hyper_parameter.epoch:
                sent_data.end_training = True
                self.__end_training = True
        super()._aggregation(sent_data=sent_data, **kwargs)

    def _stopped(self) -> bool:
        return self.__end_training or self._force_stop

    def _get_sent_data(self) -> Any:
        data = super()._get_sent_data()
        if self.__phase == Phase.STAGE_ONE:
            assert isinstance(data, ParameterMessage)
            data.parameter = self.get_block_parameter(
                parameter_dict=data.parameter, model_util=self.trainer.model_util, model_cache=self._model_cache)
            return data
        data.in_round = True
        data.other_data["check_acc"] = True
        get_logger().debug("phase 2 keys %s", data.other_data.keys())
        return data
