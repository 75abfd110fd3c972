"""Server that aggregates the workers' uploads once per round.

What a run leaves behind and when it ends follows /root/reference/simulation_lib/server/aggregation_server.py:15-184:
the initial model goes out first (``distribute_init_parameters``; warm start from ``algorithm_kwargs.global_model_path``),
every aggregated model is evaluated on the test split into ``round_record.json``, the best one is kept as
``best_global_model.pk``, the round's model is cached under ``aggregated_model/round_<n>.pk``, Shapley-style methods
ask for the metric of the initial model (``need_init_performance``, stat key 0), and ``early_stop`` ends training on
a 5-round plateau with a 0.001 threshold.  ``_get_stat_key`` stays an override point.  Not reproduced: the reference
keeps polling after it has broadcast ``end_training`` (SURVEY.md defect D8); this server stops.

The bookkeeping is split from the protocol: ``_RoundLedger`` owns the metric history, its files and the plateau
rule; ``_CheckpointStore`` owns the round-counter checkpoint (an extension: ``algorithm_kwargs.checkpoint_interval`` /
``resume_from``); ``AggregationServer`` is the protocol - collect one upload per worker, aggregate, publish.
"""
import json
import os
import pickle
from typing import Any

from ..algorithm.aggregation_algorithm import AggregationAlgorithm
from ..engine.log import get_logger
from ..engine.tensor import TensorDict, tensor_to
from ..message import FlatParameter, Message, ParameterMessage, ParameterMessageBase
from ..util.model_cache import ModelCache
from ..utils import tracing
from ..utils.observer import notify_round_end
from .server import Server

_PLATEAU_ROUNDS = 5
_PLATEAU_THRESHOLD = 0.001


def _host_copy(parameter: TensorDict) -> dict:
    return dict(tensor_to(dict(parameter), device="cpu"))


class _RoundLedger:
    """Test metrics per stat key, the files derived from them, and the early-stop rule."""

    def __init__(self, save_dir: str) -> None:
        self._save_dir = save_dir
        self.records: dict = {}
        self.best_saved = 0.0
        self.plateau_reference = 0.0
        self.plateau_length = 0

    def best_accuracy(self) -> float:
        return max(record["test_accuracy"] for record in self.records.values())

    def write(self) -> None:
        with open(os.path.join(self._save_dir, "round_record.json"), "wt", encoding="utf8") as f:
            json.dump(self.records, f)

    def add(self, key: int, metric: dict, parameter: TensorDict) -> None:
        if key in self.records:
            raise KeyError(f"round record {key} exists")
        self.records[key] = {f"test_{name}": value for name, value in metric.items()}
        self.write()
        best = self.best_accuracy()
        if best > self.best_saved:
            self.best_saved = best
            with open(os.path.join(self._save_dir, "best_global_model.pk"), "wb") as f:
                pickle.dump(_host_copy(parameter), f)

    def rename(self, old: int, new: int) -> None:
        self.records[new] = self.records.pop(old)
        self.write()

    def on_plateau(self, current_key: int) -> bool:
        """True once the best accuracy has not improved by more than the threshold for five checks in a row."""
        best = self.best_accuracy()
        if best > self.plateau_reference + _PLATEAU_THRESHOLD:
            self.plateau_reference, self.plateau_length = best, 0
            return False
        self.plateau_length += 1
        get_logger().warning("no improvement over %s for %s round(s); this round is %s below", self.plateau_reference,
                             self.plateau_length, self.plateau_reference - self.records[current_key]["test_accuracy"])
        return self.plateau_length >= _PLATEAU_ROUNDS

    # what a checkpoint must carry to continue the history
    def state(self) -> dict:
        return {"stat": self.records, "plateau": self.plateau_length, "max_acc": self.plateau_reference,
                "best_acc_saved": self.best_saved}

    def restore(self, state: dict) -> None:
        self.records = {int(key): value for key, value in state["stat"].items()}
        self.plateau_length, self.plateau_reference = state["plateau"], state["max_acc"]
        self.best_saved = state["best_acc_saved"]


class _CheckpointStore:
    """``server/checkpoint.pk``: round number, global model and ledger state, replaced atomically."""

    def __init__(self, save_dir: str, interval: int) -> None:
        self._save_dir, self.interval = save_dir, interval

    @staticmethod
    def path_in(save_dir: str) -> str:
        return os.path.join(save_dir, "server", "checkpoint.pk")

    def due(self, round_number: int) -> bool:
        return self.interval > 0 and round_number % self.interval == 0

    def save(self, round_number: int, parameter: TensorDict, ledger: _RoundLedger) -> None:
        from ..parallel import context as parallel_context

        ctx = parallel_context.current()
        if ctx is not None and ctx.rank != 0:
            return  # the replicas are identical: one writer
        target = self.path_in(self._save_dir)
        os.makedirs(os.path.dirname(target), exist_ok=True)
        scratch = target + ".tmp"
        with open(scratch, "wb") as f:
            pickle.dump({"round": round_number, "parameter": _host_copy(parameter), **ledger.state()}, f)
        os.replace(scratch, target)  # a crash mid-write leaves the previous checkpoint intact

    @classmethod
    def load(cls, location: str) -> dict:
        if os.path.isdir(location):
            location = cls.path_in(location)
        with open(location, "rb") as f:
            return pickle.load(f)


class AggregationServer(Server):
    def __init__(self, algorithm: AggregationAlgorithm, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        options = self.config.algorithm_kwargs
        self.__algorithm = algorithm
        self.__reported: set = set()
        self.__ledger = _RoundLedger(self.save_dir)
        self.__checkpoints = _CheckpointStore(self.config.save_dir, int(options.get("checkpoint_interval", 0) or 0))
        self.__early_stop = bool(options.get("early_stop", False))
        self.__finished = False
        self.__resume_model: TensorDict | None = None
        self._model_cache = ModelCache()
        self._round_number = 1
        self._compute_stat = True
        self.need_init_performance = False
        if self.__early_stop:
            get_logger().warning("training ends early on an accuracy plateau")
        if options.get("resume_from"):
            self.__continue_from(options["resume_from"])

    # ---- facts -----------------------------------------------------------------------------------------
    @property
    def early_stop(self) -> bool:
        return self.__early_stop

    @property
    def algorithm(self) -> AggregationAlgorithm:
        return self.__algorithm

    @property
    def round_number(self) -> int:
        return self._round_number

    @property
    def performance_stat(self) -> dict:
        return self.__ledger.records

    def _get_stat_key(self) -> int:
        return self._round_number

    def _stopped(self) -> bool:
        return self.__finished or self._round_number > self.config.round

    # ---- checkpoint / resume (extension over the reference) ----------------------------------------------
    @staticmethod
    def checkpoint_path(save_dir: str) -> str:
        return _CheckpointStore.path_in(save_dir)

    def __continue_from(self, location: str) -> None:
        state = _CheckpointStore.load(location)
        self.__ledger.restore(state)
        self.__resume_model = state["parameter"]
        self._round_number = int(state["round"]) + 1
        get_logger().warning("resume from %s: continue with round %d of %d", location, self._round_number,
                             self.config.round)

    # ---- first message ---------------------------------------------------------------------------------
    def _get_init_model(self) -> TensorDict:
        if self.__resume_model is not None:
            return self.__resume_model
        warm_start = self.config.algorithm_kwargs.get("global_model_path")
        if warm_start is not None:
            with open(warm_start, "rb") as f:
                return pickle.load(f)
        tester = self.tester
        tester.set_device(self._get_device(), stream=self._server_stream())
        util = tester.model_util
        if util.flat_parameter is not None:
            return FlatParameter(util.flat_parameter.detach().clone(), util.layout)
        return {name: tensor.clone() for name, tensor in util.get_parameter_dict().items()}

    def _before_start(self) -> None:
        if not self.config.distribute_init_parameters:
            if self.__resume_model is not None:
                raise ValueError("resume_from needs distribute_init_parameters")
            return
        tags: dict = {"init": True}
        if self.__resume_model is not None:
            tags["resume_round"] = self._round_number  # workers continue their own round counter
        self._send_result(ParameterMessage(in_round=True, parameter=self._get_init_model(), other_data=tags))

    def _server_exit(self) -> None:
        self.__algorithm.exit()

    # ---- collect -> aggregate -> publish -----------------------------------------------------------------
    def _process_worker_data(self, worker_id: int, data: Message | None) -> None:
        if not 0 <= worker_id < self.worker_number:
            raise IndexError(f"worker id {worker_id} out of range")
        self.__algorithm.process_worker_data(worker_id=worker_id, worker_data=data,
                                             save_dir=self.config.get_save_dir(),
                                             old_parameter_dict=self._model_cache.parameter_dict)
        self.__reported.add(worker_id)
        missing = self.worker_number - len(self.__reported)
        if missing:
            get_logger().debug("%s upload(s) in, waiting for %s more", len(self.__reported), missing)
            return
        self.__reported.clear()
        self._send_result(self._aggregate_worker_data())

    def _aggregate_worker_data(self) -> Any:
        with tracing.phase("aggregate"):
            return self.__algorithm.aggregate_worker_data()

    def _before_send_result(self, result: Message) -> None:
        if not isinstance(result, ParameterMessageBase):
            return
        if not isinstance(result, ParameterMessage):
            raise TypeError("the server publishes full models")
        is_init = "init" in result.other_data
        if is_init and self.need_init_performance and "resume_round" not in result.other_data:
            if not self.config.distribute_init_parameters:
                raise ValueError("need_init_performance needs distribute_init_parameters")
            # metric of the initial model, filed under key 0 (the Shapley engines start from it)
            self._record_compute_stat(result.parameter, keep_performance_logger=False)
            self.__ledger.rename(1, 0)
        elif not is_init and self._compute_stat:
            self._record_compute_stat(result.parameter)
            if self.early_stop and not result.end_training and self._convergent():
                result.end_training = True
        elif result.end_training:
            self._record_compute_stat(result.parameter)
        self._model_cache.cache_parameter_dict(
            result.parameter, os.path.join(self.config.save_dir, "aggregated_model", f"round_{self.round_number}.pk"))
        if not result.in_round and self.__checkpoints.due(self._round_number):
            self.__checkpoints.save(self._round_number, result.parameter, self.__ledger)
        self.__finished = self.__finished or bool(result.end_training)

    def _after_send_result(self, result: Any) -> None:
        if isinstance(result, ParameterMessageBase) and not result.in_round:
            notify_round_end(self, self._round_number)
            self._round_number += 1
        self.__algorithm.clear_worker_data()

    def _after_unicast_result(self, result: Message) -> None:
        self.__algorithm.clear_worker_data()

    # ---- evaluation bookkeeping ------------------------------------------------------------------------
    def _dump_stat(self) -> None:
        self.__ledger.write()

    def _record_compute_stat(self, parameter_dict: TensorDict, keep_performance_logger: bool = True) -> None:
        self.tester.set_visualizer_prefix(f"round: {self._round_number},")
        with tracing.phase("evaluate"):
            metric = self.get_metric(parameter_dict, keep_performance_logger=keep_performance_logger)
        self.__ledger.add(self._get_stat_key(), metric, parameter_dict)

    def _convergent(self) -> bool:
        return self.__ledger.on_plateau(self._get_stat_key())
