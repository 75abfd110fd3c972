import time


class TimeCounter:
    def __init__(self) -> None:
        self._start = time.monotonic()

    def reset_start_time(self) -> None:
        self._start = time.monotonic()

    def elapsed_milliseconds(self) -> float:
        return (time.monotonic() - self._start) * 1000
