"""Common base of servers and workers: a named role with a private copy of the config, its own output
directory and a device.

The reference's executor (/root/reference/simulation_lib/executor.py:16-96) exists to serialise co-located
roles: greenlets behind one semaphore plus a process-wide device lock, so two workers never touch the GPU at the
same time.  Here a role is a thread pinned to one GPU with its own CUDA stream, and roles that share a B200
overlap their kernels.  What is kept is the interface: ``Executor`` (``config``, ``name``, ``save_dir``,
``_get_device``, ``_release_device_lock``, ``_get_execution_context``) and ``ExecutorContext``, which now only
renames the running thread so that log lines read ``{worker 3}`` / ``{server}``.  The device lock is optional
(``device_lock=None`` is the normal case) and, when one is passed, is held from the first ``_get_device`` to
``_release_device_lock``.
"""
import copy
import os
import threading
from functools import cached_property
from typing import Any, Callable

import torch

from .engine.device import get_device


class ExecutorContext:
    """``with ExecutorContext("worker 3"):`` -- the thread carries the role's name inside the block."""

    local_data = threading.local()

    def __init__(self, name: str) -> None:
        self.__role_name = name
        self.__saved_thread_name: str | None = None

    def acquire(self) -> None:
        thread = threading.current_thread()
        self.__saved_thread_name, thread.name = thread.name, self.__role_name
        ExecutorContext.local_data.ctx = self

    def release(self) -> None:
        if self.__saved_thread_name is not None:
            threading.current_thread().name = self.__saved_thread_name
            self.__saved_thread_name = None

    def __enter__(self) -> None:
        self.acquire()

    def __exit__(self, *exc: Any) -> None:
        self.release()


class _DeviceSlot:
    """The role's GPU plus the optional exclusive hold on it."""

    def __init__(self, lock: Any, index: int | None) -> None:
        self.lock, self.index = lock, index
        self.device: torch.device | None = None
        self.holding = False
        self.peak_bytes: int | None = None   # peak allocation of the previous hold: a hint for device choice

    def acquire(self, on_first_hold: Callable | None) -> torch.device:
        if self.device is None:
            if self.lock is not None and not self.holding:
                self.lock.acquire()
                self.holding = True
                if on_first_hold is not None:
                    on_first_hold()
            self.device = get_device(max_needed_bytes=self.peak_bytes, index=self.index)
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)   # raw kernel launches of this thread go to the role's GPU
        return self.device

    def release(self) -> None:
        if not self.holding:
            return
        if self.device is not None and self.device.type == "cuda":
            stats = torch.cuda.memory_stats(device=self.device)
            if stats:
                self.peak_bytes = stats["allocated_bytes.all.peak"]
        self.lock.release()
        self.holding = False


class Executor:
    def __init__(self, config: Any, name: str, device_lock: Any = None, device_index: int | None = None) -> None:
        self.__config = copy.deepcopy(config)   # roles may edit their config (FedOBD changes epoch counts)
        self.__name = name
        self.__slot = _DeviceSlot(device_lock, device_index)

    @property
    def config(self) -> Any:
        return self.__config

    @property
    def name(self) -> str:
        return self.__name

    @cached_property
    def save_dir(self) -> str:
        root = self.config.get_save_dir()
        if not root:
            raise RuntimeError("the config has no save_dir")
        path = os.path.abspath(os.path.join(root, self.__name.replace(" ", "_")))
        os.makedirs(path, exist_ok=True)
        return path

    def _get_device(self, lock_callback: None | Callable = None) -> torch.device:
        return self.__slot.acquire(lock_callback)

    def _release_device_lock(self, **kwargs: Any) -> None:
        self.__slot.release()

    def _get_execution_context(self) -> ExecutorContext:
        return ExecutorContext(name=self.__name)

    def start(self) -> None:
        raise NotImplementedError()
