"""Process-wide observers of round boundaries (used by bench.py and the tracing hooks).
An observer is ``fn(server, round_number)`` called on the server thread right after the
round's result has been handed to the workers."""
import threading
from typing import Callable

_lock = threading.Lock()
_observers: list[Callable] = []


def add_round_observer(fn: Callable) -> None:
    with _lock:
        _observers.append(fn)


def remove_round_observer(fn: Callable) -> None:
    with _lock:
        if fn in _observers:
            _observers.remove(fn)


def notify_round_end(server, round_number: int) -> None:
    with _lock:
        observers = list(_observers)
    for fn in observers:
        fn(server, round_number)
