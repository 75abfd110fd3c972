"""Client/server endpoints: pickled objects over the topology's pipes."""


class ServerEndpoint:
    def __init__(self, topology, **kwargs) -> None:
        self._topology = topology

    @property
    def worker_num(self) -> int:
        return self._topology.worker_num

    def has_data(self, worker_id: int) -> bool:
        return self._topology.server_end(worker_id).poll()

    def get(self, worker_id: int):
        return self._topology.server_end(worker_id).recv()

    def send(self, worker_id: int, data) -> None:
        self._topology.server_end(worker_id).send(data)

    def broadcast(self, data, worker_ids=None) -> None:
        for worker_id in (range(self.worker_num) if worker_ids is None else worker_ids):
            self.send(worker_id=worker_id, data=data)

    def close(self) -> None:
        pass


class ClientEndpoint:
    def __init__(self, topology, worker_id: int, **kwargs) -> None:
        self._topology = topology
        self._worker_id = worker_id

    def has_data(self) -> bool:
        return self._topology.worker_end(self._worker_id).poll()

    def get(self):
        return self._topology.worker_end(self._worker_id).recv()

    def send(self, data) -> None:
        self._topology.worker_end(self._worker_id).send(data)

    def close(self) -> None:
        pass
