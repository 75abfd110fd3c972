"""Device enumeration / selection.  ``get_device`` picks the GPU with the most free memory,
which -- together with the reference's global device lock held until the first batch has run
(simulation_lib/executor.py:69-93, worker/worker.py:73-80) -- spreads the pool processes over
the GPUs one by one."""
import os

import torch


def get_devices() -> list:
    if torch.cuda.is_available() and os.environ.get("REF_FORCE_CPU", "0") != "1":
        n = torch.cuda.device_count()
        limit = int(os.environ.get("REF_GPUS", n))
        return [torch.device("cuda", i) for i in range(min(n, limit))]
    return [torch.device("cpu")]


def get_cpu_device():
    return torch.device("cpu")


def get_device(max_needed_bytes=None):
    devices = get_devices()
    if devices[0].type == "cpu":
        return devices[0]
    best, best_free = devices[0], -1
    for d in devices:
        free, _ = torch.cuda.mem_get_info(d)
        if free > best_free + (64 << 20):  # prefer the lowest index among (almost) equally free GPUs
            best, best_free = d, free
    return best
