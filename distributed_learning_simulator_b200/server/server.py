"""Server base: a listening loop over the endpoint, an evaluation unit, result fan-out and client sampling.

The template hooks subclasses and methods rely on are those of /root/reference/simulation_lib/server/server.py:20-134
(``tester``, ``get_metric``, ``start``, ``_before_start``, ``_server_exit``, ``_process_worker_data``,
``_before_send_result``, ``_after_send_result``, ``_send_result``, ``_select_workers``, ``_stopped``).  What sits
behind them is this framework's own:

* ``_listen`` blocks on the endpoint's condition variable until a worker it still expects has data and yields
  ``(worker_id, message)`` pairs; there is no sweep-and-sleep (reference server.py:84-85, SURVEY.md defect D7), and
  ``config.round_timeout`` turns silent workers into a ``TimeoutError``;
* ``_Evaluation`` owns the tester: created on first use on the server's device and stream, sharded over the ranks
  in multi-process mode, fed flat rows without a dict round trip when the layouts agree;
* everything the server enqueues runs on one CUDA stream of its own, so evaluation and aggregation overlap the
  workers' streams on the same GPU.
"""
import contextlib
import os
import pickle
import random
from typing import Any, Iterator

import torch

from ..engine.inferencer import Inferencer
from ..engine.log import get_logger
from ..engine.ml_type import MachineLearningPhase
from ..engine.tensor import TensorDict
from ..executor import Executor
from ..message import Message, ParameterMessage, as_flat
from ..topology.endpoint import ServerEndpoint
from ..utils import tracing


class _Evaluation:
    """The server's tester and the way models get into it."""

    def __init__(self, server: "Server") -> None:
        self._server = server
        self._tester: Inferencer | None = None

    @property
    def tester(self) -> Inferencer:
        if self._tester is None:
            self._tester = self._build()
        return self._tester

    def _build(self) -> Inferencer:
        server = self._server
        tester = server.config.create_inferencer(phase=MachineLearningPhase.Test)
        for unused in (MachineLearningPhase.Training, MachineLearningPhase.Validation):
            tester.dataset_collection.remove_dataset(phase=unused)
        tester.hook_config.summarize_executor = False
        if tester.hook_config.use_fused_kernels:
            from ..ops import gemm

            gemm.install_inference_linear(tester.model)  # evaluation GEMMs on the hand-written tcgen05 kernel
        if server._dist is not None:
            # every replica evaluates its shard of the test split; the sums are all-reduced
            tester.shard(server._dist.rank, server._dist.world, server._dist)
        return tester

    def load(self, model: TensorDict) -> None:
        tester, server = self.tester, self._server
        tester.set_device(server._get_device(), stream=server._server_stream())
        util = tester.model_util
        flat = as_flat(model)
        if flat is not None and util.flat_parameter is not None and flat.layout == util.layout:
            util.load_parameter_list(flat.flat)
        else:
            util.load_parameter_dict(model)
        util.disable_running_stats()

    def run(self, keep_performance_logger: bool) -> dict:
        tester = self.tester
        tester.hook_config.log_performance_metric = keep_performance_logger
        tester.inference()
        return tester.performance_metric.get_epoch_metrics(1)


class Server(Executor):
    # Aggregation protocols answer once every worker has reported, so a sweep takes exactly one message per
    # worker.  Per-step protocols (sign_SGD) do not: a worker whose shard holds one more batch sends two
    # messages while the others send one, and must be listened to at any time.
    _one_message_per_sweep = True

    def __init__(self, task_id: int | None, endpoint: ServerEndpoint, **kwargs: Any) -> None:
        from ..parallel import context as dist_context

        self._dist = dist_context.get_context() if dist_context.is_distributed() else None
        label = "server" if task_id is None else f"server of {task_id}"
        if self._dist is not None and self._dist.rank != 0:
            label += f" replica {self._dist.rank}"
        kwargs.setdefault("device_index", 0)
        super().__init__(name=label, **kwargs)
        self.task_id = task_id
        self._endpoint: ServerEndpoint = endpoint
        self._evaluation = _Evaluation(self)
        self._stream: Any = None
        # replicas must draw the same clients: the generator depends on the configured seed only
        self._selection_rng = random.Random(self.config.reproducible_env_config.seed + 17)

    # ---- facts -----------------------------------------------------------------------------------------
    @property
    def worker_number(self) -> int:
        return self.config.worker_number

    @property
    def tester(self) -> Inferencer:
        return self._evaluation.tester

    # ---- the server's stream ---------------------------------------------------------------------------
    def _server_stream(self) -> Any:
        device = self._get_device()
        if device.type != "cuda":
            return None
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=device)
        return self._stream

    def _stream_context(self) -> Any:
        stream = self._server_stream()
        return contextlib.nullcontext() if stream is None else torch.cuda.stream(stream)

    # ---- evaluation ------------------------------------------------------------------------------------
    def _load_into_tester(self, parameter_dict: TensorDict) -> None:
        self._evaluation.load(parameter_dict)

    def get_metric(self, parameter_dict: TensorDict | ParameterMessage, keep_performance_logger: bool = True) -> dict:
        model = parameter_dict.parameter if isinstance(parameter_dict, ParameterMessage) else parameter_dict
        self._evaluation.load(model)
        metric = self._evaluation.run(keep_performance_logger)
        self._release_device_lock()
        self.tester.offload_from_device()
        return metric

    # ---- listening loop --------------------------------------------------------------------------------
    def _listen(self) -> Iterator[tuple[int, Any]]:
        """Messages in arrival order (ties by worker id).  In one-message-per-sweep mode a worker is not listened to
        again until every other worker has been heard in the current sweep."""
        everyone = frozenset(range(self._endpoint.worker_num))
        assert len(everyone) == self.config.worker_number
        timeout = getattr(self.config, "round_timeout", None)
        expected: set = set()
        while not self._stopped():
            if not expected or not self._one_message_per_sweep:
                expected = set(everyone)
            arrived = self._endpoint.wait_for_data(expected, timeout=timeout)
            if not arrived:
                raise TimeoutError(f"no data from workers {sorted(expected)} within {timeout}s")
            for worker_id in sorted(arrived):
                expected.discard(worker_id)
                yield worker_id, self._endpoint.get(worker_id=worker_id)
                if self._stopped():
                    return

    def start(self) -> None:
        with self._get_execution_context(), self._stream_context():
            with open(os.path.join(self.save_dir, "config.pkl"), "wb") as f:
                pickle.dump(self.config, f)
            self._before_start()
            for worker_id, message in self._listen():
                get_logger().debug("message from worker %s of %s", worker_id, self._endpoint.worker_num)
                self._process_worker_data(worker_id, message)
            self._endpoint.close()
            self._server_exit()
            get_logger().debug("end server")

    # ---- template hooks --------------------------------------------------------------------------------
    def _before_start(self) -> None:
        """Runs once before the first message is awaited."""

    def _server_exit(self) -> None:
        """Runs once after the loop has ended and the endpoint is closed."""

    def _process_worker_data(self, worker_id: int, data: Message) -> None:
        raise NotImplementedError()

    def _stopped(self) -> bool:
        raise NotImplementedError()

    def _before_send_result(self, result: Message) -> None:
        """Last chance to look at / amend ``result`` before it leaves."""

    def _after_send_result(self, result: Message) -> None:
        """Runs after a broadcast answer has been handed to the endpoint."""

    def _after_unicast_result(self, result: Message) -> None:
        """Runs after per-worker answers (``other_data["worker_result"]``) have been sent."""

    # ---- answers ---------------------------------------------------------------------------------------
    def _send_result(self, result: Message) -> None:
        self._before_send_result(result=result)
        fence = getattr(self._endpoint.topology, "pre_send_fence", None)
        if fence is not None:
            fence()  # peers may still be reading this rank's upload rows
        per_worker = result.other_data.get("worker_result")
        if per_worker is not None:
            for worker_id, answer in per_worker.items():
                self._endpoint.send(worker_id=worker_id, data=answer)
            self._after_unicast_result(result=result)
            return
        chosen = self._select_workers()
        get_logger().debug("selected workers %s", chosen)
        if chosen:
            with tracing.phase("broadcast"):
                self._endpoint.broadcast(data=result, worker_ids=chosen)
        skipped = set(range(self.worker_number)) - chosen
        if skipped:
            self._endpoint.broadcast(data=None, worker_ids=skipped)  # "not selected this round"
        self._after_send_result(result=result)

    def _select_workers(self) -> set:
        population = range(self.worker_number)
        wanted = self.config.algorithm_kwargs.get("random_client_number")
        if wanted is None:
            return set(population)
        return set(self._selection_rng.sample(list(population), k=min(int(wanted), self.worker_number)))
