"""gevent stand-in on OS threads.  The reference uses greenlets only to interleave the workers
that share a process, and serialises them with a semaphore (simulation_lib/executor.py:17-38,
training.py:74), so threads + the same semaphore reproduce its scheduling: at most one
co-located worker runs at a time, the others sleep in ``gevent.sleep``."""
import threading
import time

from . import local, lock  # noqa: F401


class Greenlet(threading.Thread):
    def __init__(self, fn, *args, **kwargs) -> None:
        super().__init__(daemon=True)
        self._fn, self._args, self._kwargs = fn, args, kwargs
        self.exception = None
        self.value = None

    def run(self) -> None:
        try:
            self.value = self._fn(*self._args, **self._kwargs)
        except BaseException as e:  # noqa: BLE001
            self.exception = e


def spawn(fn, *args, **kwargs) -> Greenlet:
    g = Greenlet(fn, *args, **kwargs)
    g.start()
    return g


def joinall(greenlets, raise_error: bool = False, timeout=None):
    for g in greenlets:
        g.join(timeout)
    if raise_error:
        for g in greenlets:
            if g.exception is not None:
                raise g.exception
    return greenlets


def sleep(seconds: float = 0) -> None:
    time.sleep(seconds)
