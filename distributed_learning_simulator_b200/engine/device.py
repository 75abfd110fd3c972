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
