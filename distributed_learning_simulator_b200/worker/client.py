"""Client: a worker that talks to a server through its endpoint
(/root/reference/simulation_lib/worker/client.py:9-22).  Waiting for the server blocks on
the endpoint's condition variable instead of the reference's 0.1 s polling loop."""
from typing import Any

from ..engine.device import capture_guard
from .worker import Worker


class Client(Worker):
    def send_data_to_server(self, data: Any) -> None:
        self._endpoint.send(data)

    def _get_data_from_server(self) -> Any:
        self._release_device_lock()
        timeout = getattr(self.config, "round_timeout", None)
        # a worker that waits inside an eager section (the graph workers exchange embeddings from forward
        # hooks) must not keep the capture guard: another role's graph capture may be what everybody waits for
        with capture_guard.released():
            if not self._endpoint.wait_for_data(timeout=timeout):
                raise TimeoutError(f"{self.name}: no answer from the server within {timeout}s")
        return self._endpoint.get()
