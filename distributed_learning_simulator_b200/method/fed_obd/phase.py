"""FedOBD runs in two stages followed by a terminal state.  The reference spreads the stage logic over flags
in its server and worker classes (/root/reference/simulation_lib/method/fed_obd/phase.py:4-7 only names the
states); here the transitions are two small state machines that the roles consult."""
from dataclasses import dataclass
from enum import IntEnum, auto


class Phase(IntEnum):
    STAGE_ONE = auto()   # block-dropout uploads after every local round, sampled workers
    STAGE_TWO = auto()   # whole-model delta uploads after every epoch, everyone participates
    END = auto()


@dataclass
class ServerDecision:
    record_metric: bool        # evaluate + record the aggregated model of this step
    announce_stage_two: bool   # tell the workers to switch (rides on the broadcast)
    stop_workers: bool         # set end_training on the broadcast


class ServerPhaseMachine:
    """Stage one lasts ``stage_one_rounds`` rounds, or until the accuracy plateaus when early stopping is on;
    stage two lasts until the workers flag the end of their last epoch."""

    def __init__(self, stage_one_rounds: int, early_stop: bool) -> None:
        self.phase = Phase.STAGE_ONE
        self.stage_one_rounds = stage_one_rounds
        self.early_stop = early_stop

    def step(self, round_number: int, workers_ended: bool, metric_requested: bool, plateau: bool) -> ServerDecision:
        decision = ServerDecision(record_metric=self.phase == Phase.STAGE_ONE or metric_requested,
                                  announce_stage_two=False, stop_workers=False)
        if workers_ended:
            self.phase = Phase.END
        elif self.phase == Phase.STAGE_ONE:
            if round_number >= self.stage_one_rounds or (self.early_stop and plateau):
                self.phase = Phase.STAGE_TWO
                decision.announce_stage_two = True
        return decision

    @property
    def finished(self) -> bool:
        return self.phase == Phase.END


class WorkerPhaseMachine:
    def __init__(self) -> None:
        self.phase = Phase.STAGE_ONE
        self.last_upload_sent = False

    def observe(self, other_data: dict) -> bool:
        """Returns True exactly once: when the server's answer moves this worker to stage two."""
        if self.phase == Phase.STAGE_ONE and "phase_two" in other_data:
            self.phase = Phase.STAGE_TWO
            return True
        return False

    def closing_upload(self, epoch: int, last_epoch: int) -> bool:
        """Stage two: the upload after the last epoch closes the run."""
        if self.phase == Phase.STAGE_TWO and epoch == last_epoch:
            self.last_upload_sent = True
            return True
        return False
