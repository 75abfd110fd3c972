"""Device enumeration and placement.

Own version of ``cyy_torch_toolbox.device.{get_devices,get_device}`` (consumed at
/root/reference/simulation_lib/config.py:10,22 and executor.py:11,76-78).

B200-first placement: one process drives exactly one GPU (``LOCAL_RANK``) when launched
by ``torch.distributed.run``; a stand-alone process round-robins its workers over every
visible GPU.  180 GB of HBM3e per device makes the reference's "pick the GPU with the
most free memory" probing unnecessary, so ``get_device`` only honours an explicit
byte budget to fail loudly when a pack of workers would not fit.
"""
import os
import threading

import torch

_rr_lock = threading.Lock()
_rr_next = 0


def get_cuda_devices() -> list[torch.device]:
    if not torch.cuda.is_available():
        return []
    if "LOCAL_RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return [torch.device("cuda", int(os.environ["LOCAL_RANK"]))]
    return [torch.device("cuda", i) for i in range(torch.cuda.device_count())]


def get_devices() -> list[torch.device]:
    devices = get_cuda_devices()
    if devices and os.environ.get("DLS_FORCE_CPU", "0") != "1":
        return devices
    return [torch.device("cpu")]


def get_cpu_device() -> torch.device:
    return torch.device("cpu")


def get_device(max_needed_bytes: int | None = None, index: int | None = None) -> torch.device:
    global _rr_next
    devices = get_devices()
    if index is None:
        with _rr_lock:
            index = _rr_next
            _rr_next += 1
    device = devices[index % len(devices)]
    if device.type == "cuda" and max_needed_bytes:
        free, _ = torch.cuda.mem_get_info(device)
        if free < max_needed_bytes:
            raise RuntimeError(f"{device} has {free} free bytes, {max_needed_bytes} needed")
    return device


class CaptureGuard:
    """Keeps library-backed eager work (cuDNN/cuBLAS forward passes of evaluations and eager
    training loops, which may autotune, allocate or synchronise) away from a CUDA-graph capture
    running in another role thread of the same process.  Eager sections hold the guard shared
    (re-entrant per thread); a capture holds it exclusively and has priority over new sharers.
    Graph replays and this framework's own kernels never take it."""

    def __init__(self) -> None:
        self._cond = threading.Condition()
        self._sharers = 0
        self._capturing = False
        self._waiting_captures = 0
        self._local = threading.local()

    class _Shared:
        def __init__(self, guard: "CaptureGuard") -> None:
            self.guard = guard

        def __enter__(self) -> None:
            g = self.guard
            depth = getattr(g._local, "depth", 0)
            if depth == 0:
                with g._cond:
                    while g._capturing or g._waiting_captures:
                        g._cond.wait()
                    g._sharers += 1
            g._local.depth = depth + 1

        def __exit__(self, *exc: object) -> None:
            g = self.guard
            g._local.depth -= 1
            if g._local.depth == 0:
                with g._cond:
                    g._sharers -= 1
                    g._cond.notify_all()

    class _Exclusive:
        def __init__(self, guard: "CaptureGuard") -> None:
            self.guard = guard

        def __enter__(self) -> None:
            g = self.guard
            assert getattr(g._local, "depth", 0) == 0, "capture requested inside an eager section"
            with g._cond:
                g._waiting_captures += 1
                while g._capturing or g._sharers:
                    g._cond.wait()
                g._waiting_captures -= 1
                g._capturing = True

        def __exit__(self, *exc: object) -> None:
            g = self.guard
            with g._cond:
                g._capturing = False
                g._cond.notify_all()

    class _Released:
        """Temporarily gives up this thread's shared hold (if any) around a blocking wait for another role,
        so that a pending capture - which stops new sharers - can never be part of a wait cycle."""

        def __init__(self, guard: "CaptureGuard") -> None:
            self.guard = guard
            self.depth = 0

        def __enter__(self) -> None:
            g = self.guard
            self.depth = getattr(g._local, "depth", 0)
            if self.depth > 0:
                g._local.depth = 0
                with g._cond:
                    g._sharers -= 1
                    g._cond.notify_all()

        def __exit__(self, *exc: object) -> None:
            g = self.guard
            if self.depth > 0:
                with g._cond:
                    while g._capturing or g._waiting_captures:
                        g._cond.wait()
                    g._sharers += 1
                g._local.depth = self.depth

    def released(self) -> "CaptureGuard._Released":
        return CaptureGuard._Released(self)

    def shared(self) -> "CaptureGuard._Shared":
        return CaptureGuard._Shared(self)

    def exclusive(self) -> "CaptureGuard._Exclusive":
        return CaptureGuard._Exclusive(self)


capture_guard = CaptureGuard()


_capture_streams: dict = {}
_capture_streams_lock = threading.Lock()


def capture_stream(device: "torch.device") -> "torch.cuda.Stream":
    """ONE capture stream per device for the whole process, never destroyed.  torch's own default capture
    stream is a singleton on whichever device captured first (wrong device for the others), and a fresh stream
    per capture is worse: cuBLAS / cuDNN workspaces are cached per (handle, stream) and a workspace handed out
    during one capture comes from THAT graph's private pool -- a later stream that reuses the address would make
    another graph point into it.  Captures are serialised by ``capture_guard.exclusive()``."""
    index = device.index if device.index is not None else torch.cuda.current_device()
    with _capture_streams_lock:
        stream = _capture_streams.get(index)
        if stream is None:
            stream = _capture_streams[index] = torch.cuda.Stream(device=torch.device("cuda", index))
    return stream
