"""Lazy, optionally disk-backed value holder.

Stands in for ``cyy_naive_lib.storage.DataStorage`` as used by the reference's model
cache (/root/reference/simulation_lib/util/model_cache.py:12-50): the value lives in
memory until ``save()`` spills it to ``data_path``; reading ``.data`` after a spill
loads it back.  On a 180 GB B200 the spill is never needed, but the contract (and the
``aggregated_model/round_<n>.pk`` files it produces) is part of the session layout.
"""
import os
import pickle
from typing import Any


class DataStorage:
    def __init__(self, data: Any = None, data_path: str | None = None) -> None:
        self.__data: Any = data
        self.__data_path: str | None = data_path
        self.__synced: bool = data is None and data_path is not None

    @property
    def data_path(self) -> str | None:
        return self.__data_path

    def set_data_path(self, data_path: str) -> None:
        if data_path != self.__data_path:
            # content (if any) now belongs to a new file that has not been written yet
            if self.__data is None and self.__synced and self.__data_path is not None:
                self.__data = self.__load()
            self.__data_path = data_path
            self.__synced = False

    def set_data(self, data: Any) -> None:
        self.__data = data
        self.__synced = False

    def mark_new_data(self) -> None:
        self.__synced = False

    def has_data(self) -> bool:
        return self.__data is not None or (self.__synced and self.__data_path is not None)

    @property
    def data(self) -> Any:
        if self.__data is None and self.__synced and self.__data_path is not None:
            self.__data = self.__load()
        return self.__data

    def __load(self) -> Any:
        assert self.__data_path is not None
        with open(self.__data_path, "rb") as f:
            return pickle.load(f)

    def save(self, release: bool = False) -> None:
        """Write to ``data_path`` if dirty; ``release`` drops the in-memory copy."""
        if self.__data is None or self.__data_path is None:
            return
        if not self.__synced:
            os.makedirs(os.path.dirname(os.path.abspath(self.__data_path)), exist_ok=True)
            with open(self.__data_path, "wb") as f:
                pickle.dump(self.__data, f)
            self.__synced = True
        if release:
            self.__data = None


class TimeCounter:
    """Wall-clock stopwatch (``cyy_naive_lib.time_counter.TimeCounter``,
    used at /root/reference/simulation_lib/training.py:88,136)."""

    def __init__(self) -> None:
        import time

        self.__time = time
        self.__start = time.monotonic()

    def reset_start_time(self) -> None:
        self.__start = self.__time.monotonic()

    def elapsed_milliseconds(self) -> float:
        return (self.__time.monotonic() - self.__start) * 1000.0
