"""Worker base: owns a trainer built from its practitioner's data shard and runs one
``trainer.train()`` per round.

Template-method parity with /root/reference/simulation_lib/worker/worker.py:15-95
(name ``worker <id>[ of <task>]``, lazy ``trainer``, ``_before_training``,
``_after_training`` dumping ``hyper_parameter.pk``, ``_stopped``, ``start``; the log
prefix ``round: <n>,``).  The worker's device is fixed (``worker_id`` round-robin over the
visible GPUs) and it trains on its own CUDA stream, so the global device lock of the
reference (worker.py:73-80) is not needed for correctness.
"""
import os
from functools import cached_property
from typing import Any

import dill

from ..engine.log import get_logger
from ..engine.trainer import Trainer
from ..executor import Executor
from ..utils import tracing
from ..practitioner import Practitioner
from ..topology.endpoint import ClientEndpoint


class Worker(Executor):
    def __init__(self, task_id: int | None, endpoint: ClientEndpoint, practitioner: Practitioner,
                 **kwargs: Any) -> None:
        worker_id = practitioner.worker_id
        name = f"worker {worker_id}"
        if task_id is not None:
            name = f"worker {worker_id} of {task_id}"
        kwargs.setdefault("device_index", worker_id)
        super().__init__(name=name, **kwargs)
        self.__practitioner: Practitioner = practitioner
        self._endpoint = endpoint
        self._round_num = 0
        self._force_stop = False
        self.task_id = task_id

    @property
    def worker_id(self) -> int:
        return self.__practitioner.worker_id

    @cached_property
    def trainer(self) -> Trainer:
        return self.__practitioner.create_trainer(self.config)

    def _offload_from_device(self) -> None:
        self.trainer.offload_from_device()

    def _before_training(self) -> None:
        pass

    def _after_training(self) -> None:
        with open(os.path.join(self.save_dir, "hyper_parameter.pk"), "wb") as f:
            dill.dump(self.trainer.hyper_parameter, f)

    def _stopped(self) -> bool:
        return self._round_num > self.config.round or self._force_stop

    def start(self, **kwargs: Any) -> None:
        first_training: bool = True
        self._round_num = 1
        self._force_stop = False
        with self._get_execution_context():
            self.trainer.set_device(self._get_device())
            # every device operation of this worker (loading the global model, training,
            # building the upload) is ordered on the worker's own stream
            with self.trainer.stream_context():
                while not self._stopped():
                    if first_training:
                        self._before_training()
                        first_training = False
                        # in case worker changes round number
                        if self._stopped():
                            break
                    else:
                        self.trainer.hook_config.summarize_executor = False
                    self.trainer.disable_hook("batch_loss_logger")
                    self.trainer.set_visualizer_prefix(prefix=f"round: {self._round_num},")
                    with tracing.phase("local_train"):
                        self.trainer.train(**kwargs)
                    self._round_num += 1
                self.trainer.wait_stream()
            get_logger().debug("finish worker")
            self._endpoint.close()
            self._after_training()
            get_logger().debug("end worker")
