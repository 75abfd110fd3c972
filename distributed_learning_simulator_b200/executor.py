"""Executor base for servers and workers.

Role parity with /root/reference/simulation_lib/executor.py:16-96 (``Executor``,
``ExecutorContext``; per-executor ``save_dir``; ``_get_device``; ``_release_device_lock``).
The execution model differs on purpose: the reference runs co-located workers as
greenlets serialised by one semaphore, so they never overlap on the GPU
(executor.py:17,23-27).  Here every executor is a thread with its own CUDA stream and a
fixed device (``worker_id`` round-robin over the visible GPUs), so the workers packed on
one B200 overlap their kernels; the context only names the thread (log lines are keyed on
it) and the device lock is kept as an optional, normally uncontended, hand-shake.
"""
import copy
import os
import threading
from functools import cached_property
from typing import Any, Callable

import torch

from .engine.device import get_device


class ExecutorContext:
    local_data = threading.local()

    def __init__(self, name: str) -> None:
        self.__name: str = name
        self.__old_name: str | None = None

    def acquire(self) -> None:
        self.__old_name = threading.current_thread().name
        threading.current_thread().name = self.__name
        ExecutorContext.local_data.ctx = self

    def release(self) -> None:
        if self.__old_name is not None:
            threading.current_thread().name = self.__old_name

    def __enter__(self) -> None:
        self.acquire()

    def __exit__(self, *args: Any, **kwargs: Any) -> None:
        self.release()


class Executor:
    def __init__(self, config: Any, name: str, device_lock: Any = None, device_index: int | None = None) -> None:
        self.__config = copy.deepcopy(config)
        self.__name = name
        self.__device_lock = device_lock
        self.__hold_device_lock: bool = False
        self.__device: torch.device | None = None
        self.__device_index = device_index
        self.__used_device_memory: int | None = None

    @property
    def config(self) -> Any:
        return self.__config

    @property
    def name(self) -> str:
        return self.__name

    @cached_property
    def save_dir(self) -> str:
        assert self.config.get_save_dir()
        executor_save_dir = os.path.abspath(os.path.join(self.config.get_save_dir(), self.__name.replace(" ", "_")))
        os.makedirs(executor_save_dir, exist_ok=True)
        return executor_save_dir

    def _get_device(self, lock_callback: None | Callable = None) -> torch.device:
        if self.__device is None:
            if self.__device_lock is not None and not self.__hold_device_lock:
                self.__device_lock.acquire()
                self.__hold_device_lock = True
                if lock_callback is not None:
                    lock_callback()
            self.__device = get_device(max_needed_bytes=self.__used_device_memory, index=self.__device_index)
            if self.__device.type == "cuda":
                torch.cuda.set_device(self.__device)
        elif self.__device.type == "cuda":
            torch.cuda.set_device(self.__device)
        return self.__device

    def _get_execution_context(self) -> ExecutorContext:
        return ExecutorContext(name=self.__name)

    def _release_device_lock(self, **kwargs: Any) -> None:
        if self.__hold_device_lock:
            if self.__device is not None and self.__device.type == "cuda":
                stats = torch.cuda.memory_stats(device=self.__device)
                if stats:
                    self.__used_device_memory = stats["allocated_bytes.all.peak"]
            self.__device_lock.release()
            self.__hold_device_lock = False

    def start(self) -> None:
        raise NotImplementedError()
