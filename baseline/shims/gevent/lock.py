import threading


class BoundedSemaphore:
    def __init__(self, value: int = 1) -> None:
        self._sem = threading.BoundedSemaphore(value)

    def acquire(self, blocking: bool = True, timeout=None) -> bool:
        return self._sem.acquire(blocking, timeout) if timeout is not None else self._sem.acquire(blocking)

    def release(self) -> None:
        self._sem.release()

    def __enter__(self):
        self.acquire()
        return self

    def __exit__(self, *args) -> None:
        self.release()
