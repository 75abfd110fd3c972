"""Role wiring: which worker runs where, and how roles reach each other.

Same product as /root/reference/simulation_lib/algorithm_factory.py:12-61: a dict with the
``topology``, a ``server`` constructor and per-group ``worker`` constructors (workers are
dealt round-robin over ``parallel_number`` groups, algorithm_factory.py:38-40; explicit
practitioners are re-numbered by sorted id, :17-23; ``endpoint_kwargs["server"|"worker"]``
are passed through).  A group here is a GPU, not an OS process: all groups live in one
process and every worker is a thread with its own stream on its group's device.
"""
import functools
import itertools
from typing import Any

from .method.algorithm_factory import CentralizedAlgorithmFactory
from .topology.topology import ProcessPipeCentralTopology, ThreadCentralTopology


def create_topology(config: Any) -> Any:
    kind = getattr(config, "topology", "auto")
    from .parallel import context as dist_context

    if dist_context.is_distributed() and kind in ("auto", "device"):
        from .parallel.topology import DistributedTopology

        return DistributedTopology(worker_num=config.worker_number, ctx=dist_context.get_context())
    if kind in ("auto", "thread", "device"):
        return ThreadCentralTopology(worker_num=config.worker_number)
    if kind == "pipe":
        return ProcessPipeCentralTopology(worker_num=config.worker_number)
    raise ValueError(f"unknown topology {kind}")


def get_worker_config(config: Any, practitioners: None | set = None) -> dict:
    if practitioners is None:
        practitioners = config.create_practitioners()
    else:
        config.worker_number = len(practitioners)
        for worker_id, practitioner in enumerate(sorted(practitioners, key=lambda p: p.id)):
            assert practitioner.has_dataset(config.dc_config.dataset_name)
            practitioner.set_worker_id(worker_id)
    assert practitioners
    import distributed_learning_simulator_b200.method  # noqa: F401  (registration)

    assert CentralizedAlgorithmFactory.has_algorithm(config.distributed_algorithm), \
        f"unknown algorithm {config.distributed_algorithm}: {CentralizedAlgorithmFactory.get_algorithm_names()}"
    topology = create_topology(config)
    result: dict = {"topology": topology}
    result["server"] = {"constructor": functools.partial(
        CentralizedAlgorithmFactory.create_server,
        algorithm_name=config.distributed_algorithm,
        endpoint_kwargs=config.endpoint_kwargs.get("server", {}),
        kwargs={"config": config},
    )}
    client_config: dict = {}
    groups = max(int(config.parallel_number), 1)
    for practitioner, group in zip(sorted(practitioners, key=lambda p: p.worker_id),
                                   itertools.cycle(range(groups))):
        client_config.setdefault(group, []).append({
            "constructor": functools.partial(
                CentralizedAlgorithmFactory.create_client,
                algorithm_name=config.distributed_algorithm,
                endpoint_kwargs=config.endpoint_kwargs.get("worker", {}) | {"worker_id": practitioner.worker_id},
                kwargs={"config": config, "practitioner": practitioner},
            ),
        })
    assert client_config
    result["worker"] = client_config
    return result
