#!/usr/bin/env python3
"""Kernel timeline of graph-replayed training steps (CUPTI through torch.profiler; diagnosis only -- no
number from here is a benchmark result).  Prints, per kernel name, the launches per step, the mean
device time, and the mean gap to the previous kernel on the same stream."""
import collections
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DLS_OUTPUT_ROOT", "/tmp/dls_trace")
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from distributed_learning_simulator_b200 import load_config_from_dict, train  # noqa: E402
from distributed_learning_simulator_b200.utils.observer import add_round_observer  # noqa: E402


def main() -> None:
    rounds = 4
    prof = profile(activities=[ProfilerActivity.CUDA])
    prof.start()  # CUPTI must be started on the thread that owns it: trace the whole (short) run, cut below
    train(load_config_from_dict({
        "dataset_name": "CIFAR10", "model_name": os.environ.get("MODEL", "resnet18"), "distributed_algorithm": "fed_avg",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 1,
        "batch_size": 64, "round": rounds, "epoch": 1, "learning_rate": 0.1, "use_amp": True, "log_level": "WARNING",
        "dataset_kwargs": {"train_size": int(os.environ.get("TRAIN", "2048")), "test_size": 256},
    }))
    torch.cuda.synchronize()
    prof.stop()
    path = os.environ.get("TRACE_OUT", "gpurun_out/step_trace.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    prof.export_chrome_trace(path + ".chrome")
    events = [e for e in json.load(open(path + ".chrome"))["traceEvents"]
              if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy") and "dur" in e]
    events.sort(key=lambda e: e["ts"])
    # steady state: everything after the second-to-last aggregation kernel (= the last round)
    marks = [i for i, e in enumerate(events) if "fedavg_reduce" in e["name"]]
    if len(marks) >= 2:
        events = events[marks[-2] + 1: marks[-1] + 1]
    by_stream = collections.defaultdict(list)
    for e in events:
        by_stream[e["args"].get("stream", 0)].append(e)
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for stream, evs in by_stream.items():
        prev_end = None
        for e in evs:
            name = e["name"].replace("void ", "")[:70]
            a = agg[name]
            a[0] += 1
            a[1] += e["dur"]
            if prev_end is not None:
                a[2] += max(e["ts"] - prev_end, 0.0)
            prev_end = e["ts"] + e["dur"]
    span = events[-1]["ts"] + events[-1]["dur"] - events[0]["ts"]
    busy = sum(a[1] for a in agg.values())
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    out = {"span_us": span, "kernel_us": busy, "kernels": len(events),
           "streams": {str(k): len(v) for k, v in by_stream.items()},
           "top": [{"name": k, "count": v[0], "total_us": round(v[1], 1), "mean_us": round(v[1] / v[0], 2),
                    "mean_gap_us": round(v[2] / v[0], 2)} for k, v in rows[:45]]}
    json.dump(out, open(path, "w"), indent=1)
    os.remove(path + ".chrome")
    print(f"span {span:.0f} us, kernel time {busy:.0f} us over {len(events)} launches")
    for r in out["top"]:
        print(f"{r['total_us']:9.0f} {r['count']:6d} {r['mean_us']:8.2f} gap {r['mean_gap_us']:6.2f}  {r['name']}")


if __name__ == "__main__":
    main()
