"""Per-phase device timing (the reference has no tracing at all: only one wall-clock
``TimeCounter`` around the whole run, simulation_lib/training.py:88,136).

``phase("aggregate")`` brackets a region with CUDA events on the current stream (host clock on
CPU) and, when enabled, an NVTX range so Nsight tools show the same names.  Nothing is
synchronised while the run is going: events are resolved when ``summary()`` is called at the
end.  Enable with ``DLS_TRACE=1``; the server then writes ``<save_dir>/server/trace.json`` with
total / mean milliseconds per phase (local_train, upload, aggregate, evaluate, broadcast)."""
import contextlib
import os
import threading
import time
from typing import Any, Iterator

import torch

_lock = threading.Lock()
_records: list[tuple[str, Any, Any]] = []


def enabled() -> bool:
    return os.environ.get("DLS_TRACE", "0") == "1"


@contextlib.contextmanager
def phase(name: str) -> Iterator[None]:
    if not enabled():
        yield
        return
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.nvtx.range_push(name)
        start.record()
    else:
        start = time.perf_counter()
    try:
        yield
    finally:
        if use_cuda:
            end.record()
            torch.cuda.nvtx.range_pop()
        else:
            end = time.perf_counter()
        with _lock:
            _records.append((name, start, end))


def summary(reset: bool = True) -> dict:
    with _lock:
        records = list(_records)
        if reset:
            _records.clear()
    out: dict = {}
    for name, start, end in records:
        if isinstance(start, float):
            ms = (end - start) * 1e3
        else:
            end.synchronize()
            ms = start.elapsed_time(end)
        entry = out.setdefault(name, {"count": 0, "total_ms": 0.0, "min_ms": float("inf"), "max_ms": 0.0,
                                      "_all": []})
        entry["count"] += 1
        entry["total_ms"] += ms
        entry["min_ms"] = min(entry["min_ms"], ms)
        entry["max_ms"] = max(entry["max_ms"], ms)
        entry["_all"].append(ms)
    for entry in out.values():
        entry["mean_ms"] = entry["total_ms"] / max(entry["count"], 1)
        ordered = sorted(entry.pop("_all"))
        entry["median_ms"] = ordered[len(ordered) // 2]  # first calls include lazy initialisation / graph capture
    return out
