"""Star topologies between one server and ``worker_num`` workers.

The reference uses ``cyy_naive_lib``'s ``ProcessPipeCentralTopology`` (pickled CPU
tensors over multiprocessing pipes, polled with sleeps:
/root/reference/simulation_lib/algorithm_factory.py:26-28, server/server.py:84-85,
worker/client.py:21).  This framework ships three transports behind one interface:

* ``ThreadCentralTopology`` (default on one process): roles are threads; a hand-off is
  a reference to device memory plus a CUDA event, delivered through condition
  variables -- no polling sleeps (fixes SURVEY.md defect D7), no pickling, no D2H.
* ``ProcessPipeCentralTopology``: the reference's transport, kept for the CPU
  plumbing configuration and as an out-of-process fake for tests.
* ``parallel.DistributedTopology``: one process per GPU under ``torch.distributed``;
  control metadata is exchanged between ranks, payloads stay in peer-mapped arenas.
"""
import copy
import multiprocessing
import threading
import time
from collections import deque
from typing import Any

import torch

from ..engine.tensor import for_each_tensor, recursive_tensor_op


class TopologyClosed(RuntimeError):
    pass


class _Envelope:
    """In-process hand-off: payload + the CUDA event after which it is valid."""

    __slots__ = ("data", "event")

    def __init__(self, data: Any) -> None:
        self.data = data
        self.event = None
        if torch.cuda.is_available() and _has_cuda_tensor(data):
            self.event = torch.cuda.Event()
            self.event.record(torch.cuda.current_stream())

    def open(self) -> Any:
        if self.event is not None:
            stream = torch.cuda.current_stream()
            stream.wait_event(self.event)

            def keep(t: torch.Tensor) -> None:
                if t.is_cuda:
                    t.record_stream(stream)

            for_each_tensor(self.data, keep)
        return self.data


def _has_cuda_tensor(data: Any) -> bool:
    found = False

    def probe(t: torch.Tensor) -> None:
        nonlocal found
        found = found or t.is_cuda

    for_each_tensor(data, probe)
    return found


def _shallow_message_copy(data: Any) -> Any:
    """Each recipient gets its own message shell (fields/dicts) over shared tensors."""
    if data is None or not hasattr(data, "__dataclass_fields__"):
        return data
    new = copy.copy(data)
    for name in data.__dataclass_fields__:
        v = getattr(data, name)
        if type(v) is dict:  # plain dicts (other_data); flat / arena payloads are shared as they are
            setattr(new, name, dict(v))
    return new


class CentralTopology:
    def __init__(self, worker_num: int) -> None:
        self.worker_num = worker_num

    # server side
    def server_has_data(self, worker_id: int) -> bool:
        raise NotImplementedError

    def server_get(self, worker_id: int) -> Any:
        raise NotImplementedError

    def server_send(self, worker_id: int, data: Any) -> None:
        raise NotImplementedError

    def server_wait(self, worker_ids: set[int], timeout: float | None) -> set[int]:
        """Block until at least one of ``worker_ids`` has data; returns the ready set."""
        raise NotImplementedError

    def server_close(self) -> None:
        pass

    # worker side
    def worker_has_data(self, worker_id: int) -> bool:
        raise NotImplementedError

    def worker_get(self, worker_id: int) -> Any:
        raise NotImplementedError

    def worker_send(self, worker_id: int, data: Any) -> None:
        raise NotImplementedError

    def worker_wait(self, worker_id: int, timeout: float | None) -> bool:
        raise NotImplementedError

    def worker_close(self, worker_id: int) -> None:
        pass


class ThreadCentralTopology(CentralTopology):
    def __init__(self, worker_num: int) -> None:
        super().__init__(worker_num)
        self._to_server: list[deque] = [deque() for _ in range(worker_num)]
        self._to_worker: list[deque] = [deque() for _ in range(worker_num)]
        self._server_cv = threading.Condition()
        self._worker_cv = [threading.Condition() for _ in range(worker_num)]
        self._closed = False
        self._failure: BaseException | None = None

    def abort(self, exc: BaseException) -> None:
        """Wake every waiter with an error (a role died)."""
        self._failure = exc
        self._closed = True
        with self._server_cv:
            self._server_cv.notify_all()
        for cv in self._worker_cv:
            with cv:
                cv.notify_all()

    def _check(self) -> None:
        if self._failure is not None:
            raise TopologyClosed(f"a peer failed: {self._failure!r}")

    def server_has_data(self, worker_id: int) -> bool:
        self._check()
        return bool(self._to_server[worker_id])

    def server_get(self, worker_id: int) -> Any:
        with self._server_cv:
            while not self._to_server[worker_id]:
                self._check()
                self._server_cv.wait(1.0)
            env = self._to_server[worker_id].popleft()
        return env.open()

    def server_send(self, worker_id: int, data: Any) -> None:
        env = _Envelope(_shallow_message_copy(data))
        with self._worker_cv[worker_id]:
            self._to_worker[worker_id].append(env)
            self._worker_cv[worker_id].notify_all()

    def server_wait(self, worker_ids: set[int], timeout: float | None) -> set[int]:
        deadline = None if timeout is None else time.monotonic() + timeout
        with self._server_cv:
            while True:
                self._check()
                ready = {w for w in worker_ids if self._to_server[w]}
                if ready:
                    return ready
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    return set()
                self._server_cv.wait(min(remaining, 1.0) if remaining is not None else 1.0)

    def worker_has_data(self, worker_id: int) -> bool:
        self._check()
        return bool(self._to_worker[worker_id])

    def worker_get(self, worker_id: int) -> Any:
        cv = self._worker_cv[worker_id]
        with cv:
            while not self._to_worker[worker_id]:
                self._check()
                cv.wait(1.0)
            env = self._to_worker[worker_id].popleft()
        return env.open()

    def worker_send(self, worker_id: int, data: Any) -> None:
        env = _Envelope(data)
        with self._server_cv:
            self._to_server[worker_id].append(env)
            self._server_cv.notify_all()

    def worker_wait(self, worker_id: int, timeout: float | None) -> bool:
        deadline = None if timeout is None else time.monotonic() + timeout
        cv = self._worker_cv[worker_id]
        with cv:
            while not self._to_worker[worker_id]:
                self._check()
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    return False
                cv.wait(min(remaining, 1.0) if remaining is not None else 1.0)
        return True


def _to_cpu(data: Any) -> Any:
    return recursive_tensor_op(data, lambda t: t.detach().cpu())


class ProcessPipeCentralTopology(CentralTopology):
    """Duplex multiprocessing pipe per worker; payload tensors are moved to the CPU and
    pickled, exactly the reference's transport."""

    def __init__(self, worker_num: int, mp_context: Any = None) -> None:
        super().__init__(worker_num)
        ctx = mp_context or multiprocessing.get_context("spawn")
        self._pipes = [ctx.Pipe(duplex=True) for _ in range(worker_num)]  # (server end, worker end)

    def server_has_data(self, worker_id: int) -> bool:
        return self._pipes[worker_id][0].poll()

    def server_get(self, worker_id: int) -> Any:
        return self._pipes[worker_id][0].recv()

    def server_send(self, worker_id: int, data: Any) -> None:
        self._pipes[worker_id][0].send(_to_cpu(data))

    def server_wait(self, worker_ids: set[int], timeout: float | None) -> set[int]:
        from multiprocessing.connection import wait

        conns = {self._pipes[w][0]: w for w in worker_ids}
        ready = wait(list(conns), timeout)
        return {conns[c] for c in ready}

    def worker_has_data(self, worker_id: int) -> bool:
        return self._pipes[worker_id][1].poll()

    def worker_get(self, worker_id: int) -> Any:
        return self._pipes[worker_id][1].recv()

    def worker_send(self, worker_id: int, data: Any) -> None:
        self._pipes[worker_id][1].send(_to_cpu(data))

    def worker_wait(self, worker_id: int, timeout: float | None) -> bool:
        return self._pipes[worker_id][1].poll(timeout)

    def server_close(self) -> None:
        for s, _ in self._pipes:
            s.close()

    def worker_close(self, worker_id: int) -> None:
        self._pipes[worker_id][1].close()
