"""Client: a worker with a line to a server (role of /root/reference/simulation_lib/worker/client.py:9-22).

``send_data_to_server`` hands a message to the endpoint; ``_get_data_from_server`` sleeps on the endpoint's condition
variable until the answer is there - no 0.1 s polling loop - and gives up after ``config.round_timeout`` seconds so a
dead server becomes an error.
"""
from typing import Any

from ..engine.device import capture_guard
from .worker import Worker


class Client(Worker):
    def send_data_to_server(self, data: Any) -> None:
        self._endpoint.send(data)

    def _get_data_from_server(self) -> Any:
        self._release_device_lock()
        patience = getattr(self.config, "round_timeout", None)
        # A worker may wait from inside an eager section (the graph workers exchange embeddings from forward hooks).
        # It must not sit on the capture guard meanwhile: the capture of another role may be what everybody waits for.
        with capture_guard.released():
            answered = self._endpoint.wait_for_data(timeout=patience)
        if not answered:
            raise TimeoutError(f"{self.name}: no answer from the server within {patience}s")
        return self._endpoint.get()
