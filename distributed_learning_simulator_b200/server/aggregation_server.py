"""Server that aggregates worker uploads once per round.

Behaviour parity with /root/reference/simulation_lib/server/aggregation_server.py:15-184:
initial-model distribution (``distribute_init_parameters``, warm start from
``algorithm_kwargs.global_model_path``), per-round test metric into
``round_record.json``, ``best_global_model.pk``, cached round model
``aggregated_model/round_<n>.pk``, ``need_init_performance`` (stat key 0), early stop on a
5-round plateau with a 0.001 improvement threshold, ``_get_stat_key`` override point.
Fixed here: after broadcasting ``end_training`` the server stops (the reference keeps
polling, SURVEY.md defect D8).
"""
import json
import os
import pickle
from typing import Any

from ..algorithm.aggregation_algorithm import AggregationAlgorithm
from ..engine.log import get_logger
from ..engine.tensor import TensorDict, tensor_to
from ..message import Message, ParameterMessage, ParameterMessageBase
from ..util.model_cache import ModelCache
from ..utils import tracing
from .server import Server


class AggregationServer(Server):
    def __init__(self, algorithm: AggregationAlgorithm, **kwargs: Any) -> None:
        Server.__init__(self, **kwargs)
        self._model_cache: ModelCache = ModelCache()
        self._round_number: int = 1
        self.__worker_flag: set = set()
        self.__algorithm: AggregationAlgorithm = algorithm
        self.__stat: dict = {}
        self._compute_stat: bool = True
        self.__plateau = 0
        self.__max_acc = 0.0
        self.__best_acc_saved = 0.0
        self.__end_training_sent = False
        self.need_init_performance = False
        self.__early_stop = self.config.algorithm_kwargs.get("early_stop", False)
        if self.__early_stop:
            get_logger().warning("stop early")
        # extension over the reference (which only warm-starts from ``global_model_path``,
        # aggregation_server.py:43-52): round-counter checkpoints and resume
        self.__checkpoint_interval = int(self.config.algorithm_kwargs.get("checkpoint_interval", 0) or 0)
        self.__resumed_parameter: TensorDict | None = None
        resume_from = self.config.algorithm_kwargs.get("resume_from", None)
        if resume_from:
            self.__resume(resume_from)

    @property
    def early_stop(self) -> bool:
        return self.__early_stop

    @property
    def algorithm(self) -> AggregationAlgorithm:
        return self.__algorithm

    @property
    def round_number(self) -> int:
        return self._round_number

    # ---- checkpoint / resume (extension) ----------------------------------------------------------
    @staticmethod
    def checkpoint_path(save_dir: str) -> str:
        return os.path.join(save_dir, "server", "checkpoint.pk")

    def __resume(self, path: str) -> None:
        if os.path.isdir(path):
            path = self.checkpoint_path(path)
        with open(path, "rb") as f:
            state = pickle.load(f)
        self._round_number = int(state["round"]) + 1
        self.__stat = {int(k): v for k, v in state["stat"].items()}
        self.__plateau, self.__max_acc = state["plateau"], state["max_acc"]
        self.__best_acc_saved = state["best_acc_saved"]
        self.__resumed_parameter = state["parameter"]
        get_logger().warning("resume from %s: continue with round %d of %d", path, self._round_number,
                             self.config.round)

    def __write_checkpoint(self, parameter: TensorDict) -> None:
        from ..parallel import context as parallel_context

        ctx = parallel_context.current()
        if ctx is not None and ctx.rank != 0:
            return  # server replicas are identical: one writer is enough
        state = {"round": self._round_number, "stat": self.__stat, "plateau": self.__plateau,
                 "max_acc": self.__max_acc, "best_acc_saved": self.__best_acc_saved,
                 "parameter": dict(tensor_to(dict(parameter), device="cpu"))}
        path = self.checkpoint_path(self.config.save_dir)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path + ".tmp", "wb") as f:
            pickle.dump(state, f)
        os.replace(path + ".tmp", path)  # a crash mid-write leaves the previous checkpoint intact

    def _get_init_model(self) -> TensorDict:
        if self.__resumed_parameter is not None:
            return self.__resumed_parameter
        init_global_model_path = self.config.algorithm_kwargs.get("global_model_path", None)
        if init_global_model_path is not None:
            with open(os.path.join(init_global_model_path), "rb") as f:
                return pickle.load(f)
        tester = self.tester
        tester.set_device(self._get_device(), stream=self._server_stream())
        from ..message import FlatParameter

        flat = tester.model_util.flat_parameter
        if flat is not None:
            return FlatParameter(flat.detach().clone(), tester.model_util.layout)
        return {k: v.clone() for k, v in tester.model_util.get_parameter_dict().items()}

    def _before_start(self) -> None:
        if self.config.distribute_init_parameters:
            other_data: dict = {"init": True}
            if self.__resumed_parameter is not None:
                other_data["resume_round"] = self._round_number  # workers continue their round counter
            self._send_result(ParameterMessage(in_round=True, parameter=self._get_init_model(),
                                               other_data=other_data))
        else:
            assert self.__resumed_parameter is None, "resume_from needs distribute_init_parameters"

    def _server_exit(self) -> None:
        self.__algorithm.exit()

    def _process_worker_data(self, worker_id: int, data: Message | None) -> None:
        assert 0 <= worker_id < self.worker_number
        self.__algorithm.process_worker_data(
            worker_id=worker_id, worker_data=data, save_dir=self.config.get_save_dir(),
            old_parameter_dict=self._model_cache.parameter_dict)
        self.__worker_flag.add(worker_id)
        if len(self.__worker_flag) == self.worker_number:
            result = self._aggregate_worker_data()
            self._send_result(result)
            self.__worker_flag.clear()
        else:
            get_logger().debug("we have %s committed, and we need %s workers,skip", len(self.__worker_flag),
                               self.worker_number)

    def _aggregate_worker_data(self) -> Any:
        with tracing.phase("aggregate"):
            return self.__algorithm.aggregate_worker_data()

    def _before_send_result(self, result: Message) -> None:
        if not isinstance(result, ParameterMessageBase):
            return
        assert isinstance(result, ParameterMessage)
        if self.need_init_performance:
            assert self.config.distribute_init_parameters
        if self.need_init_performance and "init" in result.other_data and "resume_round" not in result.other_data:
            self._record_compute_stat(result.parameter, keep_performance_logger=False)
            self.__stat[0] = self.__stat.pop(1)
            self._dump_stat()
        elif self._compute_stat and "init" not in result.other_data:
            self._record_compute_stat(result.parameter)
            if not result.end_training and self.early_stop and self._convergent():
                result.end_training = True
        elif result.end_training:
            self._record_compute_stat(result.parameter)
        model_path = os.path.join(self.config.save_dir, "aggregated_model", f"round_{self.round_number}.pk")
        self._model_cache.cache_parameter_dict(result.parameter, model_path)
        if self.__checkpoint_interval > 0 and not result.in_round \
                and self._round_number % self.__checkpoint_interval == 0:
            self.__write_checkpoint(result.parameter)
        if result.end_training:
            self.__end_training_sent = True

    def _after_send_result(self, result: Any) -> None:
        if isinstance(result, ParameterMessageBase) and not result.in_round:
            from ..utils.observer import notify_round_end

            notify_round_end(self, self._round_number)
            self._round_number += 1
        self.__algorithm.clear_worker_data()

    def _after_unicast_result(self, result: Message) -> None:
        self.__algorithm.clear_worker_data()

    def _stopped(self) -> bool:
        return self._round_number > self.config.round or self.__end_training_sent

    @property
    def performance_stat(self) -> dict:
        return self.__stat

    def _get_stat_key(self) -> int:
        return self._round_number

    def _dump_stat(self) -> None:
        with open(os.path.join(self.save_dir, "round_record.json"), "wt", encoding="utf8") as f:
            json.dump(self.__stat, f)

    def _record_compute_stat(self, parameter_dict: TensorDict, keep_performance_logger: bool = True) -> None:
        self.tester.set_visualizer_prefix(f"round: {self._round_number},")
        with tracing.phase("evaluate"):
            metric = self.get_metric(parameter_dict, keep_performance_logger=keep_performance_logger)
        round_stat = {f"test_{k}": v for k, v in metric.items()}
        key = self._get_stat_key()
        assert key not in self.__stat
        self.__stat[key] = round_stat
        self._dump_stat()
        max_acc = max(t["test_accuracy"] for t in self.__stat.values())
        if max_acc > self.__best_acc_saved:
            self.__best_acc_saved = max_acc
            with open(os.path.join(self.save_dir, "best_global_model.pk"), "wb") as f:
                pickle.dump(dict(tensor_to(dict(parameter_dict), device="cpu")), f)

    def _convergent(self) -> bool:
        max_acc = max(t["test_accuracy"] for t in self.performance_stat.values())
        diff = 0.001
        if max_acc > self.__max_acc + diff:
            self.__max_acc = max_acc
            self.__plateau = 0
            return False
        get_logger().warning("max acc is %s diff is %s", self.__max_acc,
                             self.__max_acc - self.performance_stat[self._get_stat_key()]["test_accuracy"])
        self.__plateau += 1
        get_logger().warning("plateau is %s", self.__plateau)
        return self.__plateau >= 5
