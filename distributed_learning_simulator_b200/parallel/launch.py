"""``train()`` under ``torch.distributed.run``: each rank runs its local workers and a
server replica (see ``parallel/topology.py``)."""
import threading
from typing import Any

import torch.distributed as dist

from ..engine.log import add_file_handler, get_logger
from ..engine.model_util import ModelUtil
from .context import get_context, is_distributed  # noqa: F401


def _wait_or_abort(task: Any, rank: int, grace_seconds: float = 20.0) -> None:
    """Waits for the local roles.  When one of them has died, its local peers are woken by the topology, but roles
    of OTHER ranks (and local ones already inside a collective) would wait for ever: after a grace period the
    process exits non-zero, which makes the launcher (torch.distributed.run) take the whole job down."""
    import logging
    import os
    import time

    deadline = None
    while not task.wait(timeout=1.0):
        if not task.errors:
            continue
        if deadline is None:
            deadline = time.monotonic() + grace_seconds
        elif time.monotonic() > deadline:
            get_logger().error("rank %s: %s role(s) failed and the others are still blocked after %.0f s; "
                               "aborting the job: %r", rank, len(task.errors), grace_seconds, task.errors[0])
            logging.shutdown()
            os._exit(1)


def train_distributed(config: Any, practitioners: None | set, task_id: int | None) -> None:
    from ..algorithm_factory import get_worker_config
    from ..training import _Task, _last_results, _write_trace

    ctx = get_context()
    shared = [config.save_dir, config.log_file]
    dist.broadcast_object_list(shared, src=0, group=ctx.control)
    config.save_dir, config.log_file = shared
    if config.log_file and ctx.rank == 0:
        add_file_handler(config.log_file)
    config.apply_global_config()
    config.parallel_number = 1
    if ctx.use_p2p:
        dc = config.create_dataset_collection()
        util = ModelUtil(config._create_model(dc).to(ctx.device))
        util.flatten_parameters()  # same storage convention (channels-last conv weights) as every executor
        layout = util.layout
        ctx.setup_arena(layout, config.worker_number)
    worker_config = get_worker_config(config, practitioners=practitioners)
    task = _Task(task_id, config, worker_config)
    threading.current_thread().name = f"rank {ctx.rank}"
    task.start()
    _wait_or_abort(task, ctx.rank)
    task.raise_errors()
    if ctx.rank == 0:
        _write_trace(config)
    _last_results.clear()
    _last_results.update(task.results)
    _last_results["save_dir"] = config.save_dir
    dist.barrier(group=ctx.control)
    get_logger().debug("rank %s finished", ctx.rank)
    return None
