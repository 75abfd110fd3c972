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
    _one_message_per_sweep = False  # workers report per optimizer step and may be out of phase by a batch

    def __init__(self, **kwargs: Any) -> None:
        kwargs.pop("algorithm", None)
        super().__init__(**kwargs)
        self.algorithm = None
        if hasattr(self._endpoint.topology, "partial_exchange"):
            self._endpoint.topology.partial_exchange = True  # workers report at different times
        self.__gradients: dict[int, Message] = {}
        self.__epoch_reports: dict[int, Message] = {}
        self.__finished: set[int] = set()
        self.__stat: dict = {}
        self.vote_count = 0
        self.fused_vote = bool(self.config.algorithm_kwargs.get("fused_vote", True))

    @property
    def performance_stat(self) -> dict:
        return self.__stat

    def _stopped(self) -> bool:
        return len(self.__finished) == self.worker_number

    def _process_worker_data(self, worker_id: int, data: Message) -> None:
        """Workers step in lock-step, but shards need not hold the same number of batches: a worker
        that has finished its epoch waits for the epoch acknowledgement while the others keep
        voting among themselves."""
        if data.end_training:
            self.__finished.add(worker_id)
        elif "gradient" in data.other_data:
            self.__gradients[worker_id] = data
        else:
            self.__epoch_reports[worker_id] = data
        active = set(range(self.worker_number)) - self.__finished
        stepping = active - set(self.__epoch_reports)
        if stepping and all(w in self.__gradients for w in stepping):
            gradients, self.__gradients = self.__gradients, {}
            n = int(next(iter(gradients.values())).other_data["numel"])
            rows = [gradients[w].other_data["gradient"] for w in sorted(gradients)]
            self.vote_count += 1
            if self.fused_vote:
                # no vote tensor: every worker runs the majority over the packed rows inside its own update
                # kernel (rows of other ranks are peer-mapped words, see parallel/topology.py)
                answer: Any = {"sign_rows": rows, "numel": n}
            else:
                answer = misc_ops.sign_majority([misc_ops.as_local_words(r) for r in rows], n)
            self._endpoint.broadcast(Message(in_round=True, other_data={"gradient": answer}), worker_ids=set(gradients))
        if active and all(w in self.__epoch_reports for w in active):
            reports, self.__epoch_reports = self.__epoch_reports, {}
            epoch = next(iter(reports.values())).other_data["epoch"]
            for msg in reports.values():
                if isinstance(msg, ParameterMessage):
                    self.tester.set_visualizer_prefix(f"epoch: {epoch},")
                    metric = self.get_metric(msg.parameter)
                    self.__stat[epoch] = {f"test_{k}": v for k, v in metric.items()}
                    get_logger().info("epoch: %s, test loss: %e, test accuracy: %.2f%%", epoch, metric["loss"],
                                      metric["accuracy"] * 100)
                    with open(os.path.join(self.save_dir, "round_record.json"), "wt", encoding="utf8") as f:
                        json.dump(self.__stat, f)
            self._endpoint.broadcast(Message(in_round=True, other_data={"epoch": epoch}), worker_ids=set(reports))
