"""Launcher and concurrent-task manager.

API parity with /root/reference/simulation_lib/training.py:1-169: ``train(config,
practitioners=None)`` blocks and logs ``training use %s seconds`` when no practitioners
are given, otherwise returns a ``task_id`` immediately; ``get_training_result(task_id,
timeout)`` returns ``{"performance": ..., "sv": ...}`` with Shapley values re-keyed from
``worker_id`` to ``practitioner_id`` (training.py:140-169).

The reference spawns one OS process per GPU plus a server and a manager process and runs
co-located workers as greenlets (training.py:103-125, 47-75).  Here a task is a set of
threads inside one process: one per worker (each with its own CUDA stream on its GPU) and
one for the server; payloads move by reference in HBM.  A role that dies aborts the
task's topology so every peer wakes up with an error instead of hanging.
"""
import copy
import os
import threading
import uuid
from typing import Any

# save memory/latency when many workers are packed: load kernels lazily (reference: training.py:6)
os.environ.setdefault("CUDA_MODULE_LOADING", "LAZY")

from .algorithm_factory import get_worker_config  # noqa: E402
from .engine.log import add_file_handler, get_logger  # noqa: E402
from .engine.storage import TimeCounter  # noqa: E402

tasks: dict = {}
task_results: dict = {}


class _Task:
    def __init__(self, task_id: int | None, config: Any, worker_config: dict) -> None:
        self.task_id = task_id
        self.config = config
        self.topology = worker_config["topology"]
        self.threads: list[threading.Thread] = []
        self.results: dict = {}
        self.errors: list[BaseException] = []
        self.server: Any = None
        self.workers: list = []
        self._lock = threading.Lock()
        extra = {"task_id": task_id, "device_lock": None}
        extra_ep = {"topology": self.topology}
        server_config = worker_config.get("server")
        if server_config is not None:
            self.server = server_config["constructor"](extra_kwargs=dict(extra), extra_endpoint_kwargs=dict(extra_ep))
        is_local = getattr(self.topology, "is_local", lambda worker_id: True)
        for group in worker_config["worker"].values():
            for wc in group:
                worker_id = wc["constructor"].keywords["endpoint_kwargs"]["worker_id"]
                if is_local(worker_id):
                    self.workers.append(wc["constructor"](extra_kwargs=dict(extra),
                                                          extra_endpoint_kwargs=dict(extra_ep)))

    def _run(self, role: Any, is_server: bool) -> None:
        try:
            role.start()
            if is_server:
                res: dict = {}
                if hasattr(role.algorithm, "shapley_values"):
                    res["sv"] = role.algorithm.shapley_values
                res |= {"performance": role.performance_stat}
                with self._lock:
                    self.results |= res
                get_logger().info("stop server")
        except BaseException as e:  # noqa: BLE001 -- must wake the peers whatever happened
            get_logger().exception("%s failed", getattr(role, "name", role))
            with self._lock:
                self.errors.append(e)
            if hasattr(self.topology, "abort"):
                self.topology.abort(e)

    def start(self) -> None:
        if self.server is not None:
            self.threads.append(threading.Thread(target=self._run, args=(self.server, True), daemon=True))
        for w in self.workers:
            self.threads.append(threading.Thread(target=self._run, args=(w, False), daemon=True))
        for t in self.threads:
            t.start()

    def wait(self, timeout: float | None = None) -> bool:
        """True when every role has finished."""
        import time

        deadline = None if timeout is None else time.monotonic() + timeout
        for t in self.threads:
            remaining = None if deadline is None else max(deadline - time.monotonic(), 0.0)
            t.join(remaining)
            if t.is_alive():
                return False
        return True

    def raise_errors(self) -> None:
        if self.errors:
            raise RuntimeError(f"task {self.task_id} failed") from self.errors[0]


def train(config: Any, practitioners: None | set = None) -> int | None:
    # we need to deepcopy config for concurrent runs
    config = copy.deepcopy(config)
    timer = TimeCounter()
    task_id = None
    handler = None
    if practitioners is None:
        if config.log_file:
            handler = add_file_handler(config.log_file)
    else:
        task_id = uuid.uuid4().int
    from .parallel import launch

    if launch.is_distributed():
        result = launch.train_distributed(config, practitioners, task_id)
        get_logger().info("training use %s seconds", timer.elapsed_milliseconds() / 1000)
        return result
    worker_config = get_worker_config(config, practitioners=practitioners)
    assert worker_config["topology"].worker_num == config.worker_number
    task = _Task(task_id, config, worker_config)
    task.start()
    if practitioners is not None:
        tasks[task_id] = {
            "task": task,
            "practitioner_ids": {practitioner.id for practitioner in practitioners},
            "config": config,
        }
        task_results[task_id] = {}
        return task_id
    task.wait()
    task.raise_errors()
    from .parallel import context as dist_context

    if dist_context.current() is not None:
        dist_context.current().finish_reductions()  # verdict of the last fused cross-GPU reduction
    _write_trace(config)
    _last_results.clear()
    _last_results.update(task.results)
    _last_results["save_dir"] = config.save_dir
    get_logger().info("training use %s seconds", timer.elapsed_milliseconds() / 1000)
    return None


_last_results: dict = {}


def _write_trace(config: Any) -> None:
    """Per-phase device timings of the run (DLS_TRACE=1), written once every role has finished."""
    from .utils import tracing

    if not tracing.enabled():
        return
    import json

    path = os.path.join(config.save_dir, "server")
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "trace.json"), "wt", encoding="utf8") as f:
        json.dump(tracing.summary(), f, indent=1)


def get_last_training_result() -> dict:
    """Result dict of the most recent blocking ``train()`` call in this process."""
    return dict(_last_results)


def get_training_result(task_id: int, timeout: None | float = None) -> None | dict:
    entry = tasks[task_id]
    task: _Task = entry["task"]
    if not task.wait(timeout):
        return None
    tasks.pop(task_id)
    task.raise_errors()
    task_results[task_id] |= task.results
    get_logger().info("finish task %s", task_id)
    stats: dict = {}
    practitioner_ids = entry["practitioner_ids"]
    config = entry["config"]
    assert practitioner_ids is not None
    for k, v in task_results[task_id].items():
        if k != "sv":
            stats[k] = v
            continue
        sv_dict: dict = {}
        for round_number, tmp_sv_dict in v.items():
            sv_dict[round_number] = {}
            for practitioner_id, worker_id in zip(sorted(practitioner_ids), range(config.worker_number)):
                sv_dict[round_number][practitioner_id] = tmp_sv_dict[worker_id]
        stats[k] = sv_dict
    task_results.pop(task_id)
    return stats
