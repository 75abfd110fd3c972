"""Worker base: a trainer over the practitioner's data shard, one ``trainer.train()`` per federated round.

Public surface of /root/reference/simulation_lib/worker/worker.py:15-95 (name ``worker <id>[ of <task>]``, lazy
``trainer``, the ``_before_training`` / ``_after_training`` / ``_stopped`` hooks, ``start``, the ``round: <n>,`` log
prefix, ``hyper_parameter.pk``).  The execution model is this framework's own: a worker is pinned to one GPU
(``worker_id`` round-robin over the visible devices) and owns a CUDA stream on it, so co-located workers overlap
and the reference's global device lock (worker.py:73-80) has no counterpart.
"""
import os
from functools import cached_property
from typing import Any, Iterator

import dill

from ..engine.log import get_logger
from ..engine.trainer import Trainer
from ..executor import Executor
from ..practitioner import Practitioner
from ..topology.endpoint import ClientEndpoint
from ..utils import tracing


class Worker(Executor):
    def __init__(self, task_id: int | None, endpoint: ClientEndpoint, practitioner: Practitioner,
                 **kwargs: Any) -> None:
        self.__practitioner = practitioner
        self.task_id = task_id
        label = f"worker {practitioner.worker_id}" + ("" if task_id is None else f" of {task_id}")
        kwargs.setdefault("device_index", practitioner.worker_id)
        super().__init__(name=label, **kwargs)
        self._endpoint = endpoint
        self._round_num = 0
        self._force_stop = False

    # ---- identity / resources ------------------------------------------------------------------------
    @property
    def worker_id(self) -> int:
        return self.__practitioner.worker_id

    @cached_property
    def trainer(self) -> Trainer:
        return self.__practitioner.create_trainer(self.config)

    def _offload_from_device(self) -> None:
        self.trainer.offload_from_device()

    # ---- template hooks ------------------------------------------------------------------------------
    def _before_training(self) -> None:
        """Runs once, on the worker's stream, before the first round."""

    def _after_training(self) -> None:
        with open(os.path.join(self.save_dir, "hyper_parameter.pk"), "wb") as f:
            dill.dump(self.trainer.hyper_parameter, f)

    def _stopped(self) -> bool:
        return self._force_stop or self._round_num > self.config.round

    # ---- round loop ----------------------------------------------------------------------------------
    def _rounds(self) -> Iterator[int]:
        """Yields the number of every round to train.  ``_before_training`` may itself end the run or move the
        round counter (a resumed server, an unselected worker), hence the re-check after it."""
        self._round_num, self._force_stop = 1, False
        prepared = False
        while not self._stopped():
            if not prepared:
                self._before_training()
                prepared = True
                if self._stopped():
                    return
            yield self._round_num
            self._round_num += 1

    def _train_round(self, round_number: int, first: bool, **kwargs: Any) -> None:
        trainer = self.trainer
        if not first:
            trainer.hook_config.summarize_executor = False
        trainer.disable_hook("batch_loss_logger")
        trainer.set_visualizer_prefix(prefix=f"round: {round_number},")
        with tracing.phase("local_train"):
            trainer.train(**kwargs)

    def start(self, **kwargs: Any) -> None:
        with self._get_execution_context():
            self.trainer.set_device(self._get_device())
            # loading the global model, training and building the upload are all ordered on this stream
            with self.trainer.stream_context():
                first = True
                for round_number in self._rounds():
                    self._train_round(round_number, first, **kwargs)
                    first = False
                self.trainer.wait_stream()
            get_logger().debug("finish worker")
            self._endpoint.close()
            self._after_training()
            get_logger().debug("end worker")
