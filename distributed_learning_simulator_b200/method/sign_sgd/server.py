"""signSGD server: element-wise majority vote over the workers' packed sign words
(``ops.misc.sign_majority``: popcount across workers, emits -1/0/+1) once per optimizer
step; per-epoch evaluation of worker 0's model into ``round_record.json``."""
import json
import os
from typing import Any

from ...engine.log import get_logger
from ...message import Message, ParameterMessage
from ...ops import misc as misc_ops
from ...server.server import Server


class SignSGDServer(Server):
    def __init__(self, **kwargs: Any) -> None:
        kwargs.pop("algorithm", None)
        super().__init__(**kwargs)
        self.algorithm = None
        self.__pending: dict[int, Message] = {}
        self.__finished: set[int] = set()
        self.__stat: dict = {}
        self.vote_count = 0

    @property
    def performance_stat(self) -> dict:
        return self.__stat

    def _stopped(self) -> bool:
        return len(self.__finished) == self.worker_number

    def _process_worker_data(self, worker_id: int, data: Message) -> None:
        if data.end_training:
            self.__finished.add(worker_id)
            return
        self.__pending[worker_id] = data
        if len(self.__pending) < self.worker_number - len(self.__finished):
            return
        pending, self.__pending = self.__pending, {}
        first = next(iter(pending.values()))
        if "gradient" in first.other_data:
            n = int(first.other_data["numel"])
            rows = [pending[w].other_data["gradient"] for w in sorted(pending)]
            vote = misc_ops.sign_majority(rows, n)
            self.vote_count += 1
            self._endpoint.broadcast(Message(in_round=True, other_data={"gradient": vote}), worker_ids=set(pending))
            return
        epoch = first.other_data["epoch"]
        for msg in pending.values():
            if isinstance(msg, ParameterMessage):
                self.tester.set_visualizer_prefix(f"epoch: {epoch},")
                metric = self.get_metric(msg.parameter)
                self.__stat[epoch] = {f"test_{k}": v for k, v in metric.items()}
                get_logger().info("epoch: %s, test loss: %e, test accuracy: %.2f%%", epoch, metric["loss"],
                                  metric["accuracy"] * 100)
                with open(os.path.join(self.save_dir, "round_record.json"), "wt", encoding="utf8") as f:
                    json.dump(self.__stat, f)
        self._endpoint.broadcast(Message(in_round=True, other_data={"epoch": epoch}), worker_ids=set(pending))
