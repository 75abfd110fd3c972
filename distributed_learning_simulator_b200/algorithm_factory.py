"""Role wiring: which worker runs in which group, and how roles reach each other.

The product has the shape of /root/reference/simulation_lib/algorithm_factory.py:12-61 because ``training._Task``
and user code consume it: ``{"topology": ..., "server": {"constructor": partial}, "worker": {group: [{"constructor":
partial}, ...]}}``.  Workers are dealt round-robin over ``parallel_number`` groups (:38-40), explicit practitioners
are re-numbered by sorted id (:17-23), ``endpoint_kwargs["server" | "worker"]`` are passed through.  A group here is
a GPU, not an OS process: all groups live in one process and every worker is a thread with its own stream on its
group's device; under ``torch.distributed.run`` the topology spans the ranks.
"""
from functools import partial
from typing import Any, Iterable

from .method.algorithm_factory import CentralizedAlgorithmFactory
from .topology.topology import ProcessPipeCentralTopology, ThreadCentralTopology


def create_topology(config: Any) -> Any:
    from .parallel import context as dist_context

    kind = getattr(config, "topology", "auto")
    workers = config.worker_number
    if kind in ("auto", "device") and dist_context.is_distributed():
        from .parallel.topology import DistributedTopology

        return DistributedTopology(worker_num=workers, ctx=dist_context.get_context())
    in_process = {"auto": ThreadCentralTopology, "thread": ThreadCentralTopology, "device": ThreadCentralTopology,
                  "pipe": ProcessPipeCentralTopology}
    if kind not in in_process:
        raise ValueError(f"unknown topology {kind}")
    return in_process[kind](worker_num=workers)


def _adopt_practitioners(config: Any, practitioners: Iterable) -> list:
    """Caller-supplied practitioners define the worker set: ids 0..n-1 in the order of their practitioner ids."""
    ordered = sorted(practitioners, key=lambda p: p.id)
    dataset = config.dc_config.dataset_name
    for worker_id, practitioner in enumerate(ordered):
        if not practitioner.has_dataset(dataset):
            raise ValueError(f"practitioner {practitioner.id} holds no {dataset} data")
        practitioner.set_worker_id(worker_id)
    config.worker_number = len(ordered)
    return ordered


def _check_algorithm(name: str) -> None:
    import distributed_learning_simulator_b200.method  # noqa: F401  (methods register themselves on import)

    if not CentralizedAlgorithmFactory.has_algorithm(name):
        raise KeyError(f"unknown algorithm {name}: {CentralizedAlgorithmFactory.get_algorithm_names()}")


def get_worker_config(config: Any, practitioners: None | set = None) -> dict:
    members = list(config.create_practitioners()) if practitioners is None else _adopt_practitioners(config, practitioners)
    if not members:
        raise ValueError("no practitioners")
    algorithm = config.distributed_algorithm
    _check_algorithm(algorithm)
    server = partial(CentralizedAlgorithmFactory.create_server, algorithm_name=algorithm,
                     endpoint_kwargs=config.endpoint_kwargs.get("server", {}), kwargs={"config": config})
    group_count = max(int(config.parallel_number), 1)
    groups: dict[int, list] = {}
    for position, practitioner in enumerate(sorted(members, key=lambda p: p.worker_id)):
        endpoint_kwargs = {**config.endpoint_kwargs.get("worker", {}), "worker_id": practitioner.worker_id}
        client = partial(CentralizedAlgorithmFactory.create_client, algorithm_name=algorithm,
                         endpoint_kwargs=endpoint_kwargs, kwargs={"config": config, "practitioner": practitioner})
        groups.setdefault(position % group_count, []).append({"constructor": client})
    return {"topology": create_topology(config), "server": {"constructor": server}, "worker": groups}
