"""FedOBD server.  Same observable behaviour as the reference's (/root/reference/simulation_lib/method/fed_obd/
server.py:10-61: quantized broadcasts, a switch to stage two after ``config.round`` rounds or on a plateau,
the run ends when the workers flag ``end_training``), driven by ``ServerPhaseMachine``."""
from typing import Any

from ...engine.log import get_logger
from ..common_import import AggregationServer, FedAVGAlgorithm, ParameterMessageBase, QuantServerEndpoint
from .phase import Phase, ServerPhaseMachine


class FedOBDServer(AggregationServer):
    def __init__(self, **kwargs: Any) -> None:
        kwargs.pop("algorithm", None)
        super().__init__(algorithm=FedAVGAlgorithm(), **kwargs)
        if not isinstance(self._endpoint, QuantServerEndpoint):
            raise TypeError("FedOBD needs a quantizing server endpoint")
        self._endpoint.quant_broadcast = True
        self.__machine = ServerPhaseMachine(stage_one_rounds=self.config.round, early_stop=self.early_stop)

    @property
    def phase(self) -> Phase:
        return self.__machine.phase

    def _stopped(self) -> bool:
        return self.__machine.finished

    def _select_workers(self) -> set:
        # client sampling is a stage-one device; afterwards every worker trains every epoch
        if self.__machine.phase == Phase.STAGE_ONE:
            return super()._select_workers()
        return set(range(self.worker_number))

    def _get_stat_key(self) -> int:
        # stage two records one entry per epoch, so the keys simply keep counting
        stat = self.performance_stat
        return super()._get_stat_key() if not stat else max(stat) + 1

    def _aggregate_worker_data(self) -> ParameterMessageBase:
        result: ParameterMessageBase = super()._aggregate_worker_data()
        before = self.__machine.phase
        decision = self.__machine.step(
            round_number=self.round_number, workers_ended=bool(result.end_training),
            metric_requested="check_acc" in result.other_data,
            plateau=before == Phase.STAGE_ONE and self.early_stop and self._convergent())
        self._compute_stat = decision.record_metric
        if decision.announce_stage_two:
            get_logger().warning("switch to phase 2")
            result.other_data["phase_two"] = True
        if decision.stop_workers:
            result.end_training = True
        return result
