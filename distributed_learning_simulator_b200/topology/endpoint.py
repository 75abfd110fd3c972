"""Client/server endpoints over a star topology.

Same surface as ``cyy_naive_lib.topology.cs_endpoint`` as consumed by the reference
(/root/reference/simulation_lib/server/server.py:66-78,88,111,117-120,
worker/client.py:11,18,22, worker/worker.py:92): ``send/get/has_data/broadcast/close``,
plus event-driven ``wait_*`` so callers never sleep-poll.
"""
from typing import Any, Iterable

from .topology import CentralTopology


class ServerEndpoint:
    def __init__(self, topology: CentralTopology, **kwargs: Any) -> None:
        self._topology = topology

    @property
    def topology(self) -> CentralTopology:
        return self._topology

    @property
    def worker_num(self) -> int:
        return self._topology.worker_num

    def has_data(self, worker_id: int) -> bool:
        return self._topology.server_has_data(worker_id)

    def get(self, worker_id: int) -> Any:
        return self._topology.server_get(worker_id)

    def send(self, worker_id: int, data: Any) -> None:
        self._topology.server_send(worker_id, data)

    def broadcast(self, data: Any, worker_ids: Iterable[int] | None = None) -> None:
        if worker_ids is None:
            worker_ids = range(self.worker_num)
        for worker_id in sorted(worker_ids):
            self.send(worker_id=worker_id, data=data)

    def wait_for_data(self, worker_ids: set[int], timeout: float | None = None) -> set[int]:
        return self._topology.server_wait(worker_ids, timeout)

    def close(self) -> None:
        self._topology.server_close()


class ClientEndpoint:
    def __init__(self, topology: CentralTopology, worker_id: int, **kwargs: Any) -> None:
        self._topology = topology
        self._worker_id = worker_id

    @property
    def worker_id(self) -> int:
        return self._worker_id

    def has_data(self) -> bool:
        return self._topology.worker_has_data(self._worker_id)

    def get(self) -> Any:
        return self._topology.worker_get(self._worker_id)

    def send(self, data: Any) -> None:
        self._topology.worker_send(self._worker_id, data)

    def wait_for_data(self, timeout: float | None = None) -> bool:
        return self._topology.worker_wait(self._worker_id, timeout)

    def close(self) -> None:
        self._topology.worker_close(self._worker_id)
