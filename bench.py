#!/usr/bin/env python3
"""Headline benchmark: federated rounds of fed_avg ResNet-18 on synthetic CIFAR-10-shape
data, one worker per B200 (BASELINE.json config 2).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A "step" is one federated round: every worker trains one local epoch (batch 64, SGD with
cosine schedule, bf16 autocast, best-model-by-validation) on its 6 250-sample shard,
uploads its delta, the server aggregates + broadcasts and evaluates the new global model
on the test split.  ``value`` is worker samples/s summed over all GPUs (training samples
consumed per second of the timed rounds); rounds/s is reported alongside.  Weak scaling:
per-GPU work is fixed (train 6 250 / test 1 250 samples per worker; N=8 is exactly the
50 000/10 000 CIFAR-10 shape).

Two measurements per run, each with W untimed + K timed rounds bracketed by
barrier + synchronize and timed with CUDA events (max over ranks):
  * device-resident data  -> ``value``
  * end to end            -> ``e2e``: every round the shard is copied host->device from
    pinned memory and the round's metrics are read back device->host.
``--impl reference`` runs the unmodified reference (baseline/_ref) on the same config.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TRAIN_PER_WORKER = 6250
TEST_PER_WORKER = 1250
BATCH = 64


def bench_config(n_workers: int, rounds: int, storage: str):
    from distributed_learning_simulator_b200 import load_config_from_dict

    return load_config_from_dict({
        "dataset_name": "CIFAR10", "model_name": "resnet18", "distributed_algorithm": "fed_avg",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR",
        "worker_number": n_workers, "batch_size": BATCH, "round": rounds, "epoch": 1, "learning_rate": 0.1,
        "use_amp": True, "log_level": "WARNING",
        "dataset_kwargs": {"train_size": TRAIN_PER_WORKER * n_workers, "test_size": TEST_PER_WORKER * n_workers},
        "dataloader_kwargs": {"storage": storage},
    })


class RoundTimer:
    """Brackets rounds (warmup, warmup+steps] with barrier + synchronize + CUDA events."""

    def __init__(self, warmup: int, steps: int, sampler=None) -> None:
        import torch

        self.torch = torch
        self.warmup, self.steps = warmup, steps
        self.start_evt = torch.cuda.Event(enable_timing=True)
        self.end_evt = torch.cuda.Event(enable_timing=True)
        self.wall = [0.0, 0.0]
        self.launches = [0, 0]
        self.sampler = sampler
        self.done = threading.Event()

    def _fence(self) -> None:
        torch = self.torch
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def __call__(self, server, round_number: int) -> None:
        from distributed_learning_simulator_b200.ops import runtime

        if round_number == self.warmup:
            self._fence()
            if self.sampler is not None:
                self.sampler.start()
            self.launches[0] = runtime.launch_count()
            self.wall[0] = time.perf_counter()
            self.start_evt.record()
        elif round_number == self.warmup + self.steps:
            self._fence()
            self.end_evt.record()
            self.end_evt.synchronize()
            self.wall[1] = time.perf_counter()
            self.launches[1] = runtime.launch_count()
            self.done.set()

    def elapsed_ms(self) -> float:
        return float(self.start_evt.elapsed_time(self.end_evt))


def run_ours(args) -> dict:
    import torch

    from distributed_learning_simulator_b200 import train
    from distributed_learning_simulator_b200.engine import graph_step
    from distributed_learning_simulator_b200.engine.metric import PerformanceMetric
    from distributed_learning_simulator_b200.ops import runtime
    from distributed_learning_simulator_b200.utils.clocks import ClockSampler
    from distributed_learning_simulator_b200.utils.observer import add_round_observer, remove_round_observer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1, f"launched with WORLD_SIZE={world} but --gpus {args.gpus}"
    n_workers = args.gpus
    if world == 1 and args.gpus > 1:
        assert torch.cuda.device_count() >= args.gpus
    torch.cuda.set_device(local_rank)
    runtime.require()
    rounds = args.warmup + args.steps
    results = {}
    for mode, storage in (("device", "device"), ("e2e", "pinned_host")):
        sampler = ClockSampler(gpu_index=local_rank) if (mode == "device" and rank == 0) else None
        timer = RoundTimer(args.warmup, args.steps, sampler)
        add_round_observer(timer)
        PerformanceMetric.d2h_bytes = 0
        counters0 = (runtime.launch_count(), graph_step.CAPTURE_LAUNCHES[0], graph_step.OWN_REPLAYED_KERNELS[0])
        config = bench_config(n_workers, rounds, storage)
        try:
            train(config)
        finally:
            remove_round_observer(timer)
        # every round launches the same kernels, so the timed rounds' share of the run's exact totals
        # (taken after all role threads have finished) is exact; warm-up/capture launches are excluded
        direct = (runtime.launch_count() - counters0[0]) - (graph_step.CAPTURE_LAUNCHES[0] - counters0[1])
        replayed = graph_step.OWN_REPLAYED_KERNELS[0] - counters0[2]
        launches = (direct + replayed) * args.steps / rounds
        clocks = sampler.stop() if sampler is not None else None
        assert timer.done.is_set(), "the timed rounds did not complete"
        ms = timer.elapsed_ms()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        h2d = graph_step.H2D_BYTES[0]
        d2h = PerformanceMetric.d2h_bytes
        if world > 1:
            t = torch.tensor([h2d, d2h, launches], dtype=torch.float64, device="cuda")
            torch.distributed.all_reduce(t)  # whole-job totals
            h2d, d2h, launches = float(t[0].item()), float(t[1].item()), float(t[2].item())
        samples = TRAIN_PER_WORKER * n_workers * args.steps
        results[mode] = {
            "ms": ms, "samples_per_s": samples / (ms / 1e3), "rounds_per_s": args.steps / (ms / 1e3),
            "wall_ms": (timer.wall[1] - timer.wall[0]) * 1e3, "clocks": clocks,
            "launches": launches,
            "h2d_bytes_per_step": h2d / max(rounds, 1) if storage == "pinned_host" else 0,
            "d2h_bytes_per_step": d2h / max(rounds, 1),
        }
        graph_step.H2D_BYTES[0] = 0
    dev, e2e = results["device"], results["e2e"]
    return {
        "metric": "worker samples/sec (fed_avg ResNet-18, CIFAR-10 shape, federated rounds)",
        "value": dev["samples_per_s"], "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev["ms"] / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ours",
        "rounds_per_s": dev["rounds_per_s"],
        "config": {"model": "resnet18 (CIFAR stem, 11.17M params)", "algorithm": "fed_avg",
                   "workers": n_workers, "global_batch": BATCH * n_workers, "seq_len": None,
                   "local_batch": BATCH, "local_epochs_per_round": 1,
                   "train_samples_per_worker": TRAIN_PER_WORKER, "test_samples": TEST_PER_WORKER * n_workers,
                   "parallelism": f"fl-dp{args.gpus} (one worker per GPU)",
                   "l2_policy": "inputs larger than L2: each round streams a 77 MB shard + 45 MB model/optimizer "
                                "state per worker; no explicit flush"},
        "clocks": dev["clocks"],
        "e2e": {"value": e2e["samples_per_s"], "unit": "samples/s", "ms_per_step": e2e["ms"] / args.steps,
                "h2d_bytes_per_step": e2e["h2d_bytes_per_step"], "d2h_bytes_per_step": e2e["d2h_bytes_per_step"]},
        "gpu_launches": int(dev["launches"]),
        "wall_ms_per_step": dev["wall_ms"] / args.steps,
    }


def run_reference(args) -> dict:
    from baseline.install_ref import ReferenceUnavailable, ensure_installed

    try:
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            ensure_installed()
    except ReferenceUnavailable as e:
        return {"impl": "reference", "unavailable": str(e)[:300].replace("\n", " "), "exit_code": 1}
    try:
        from baseline.reference_arm import run as run_ref
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference harness not importable: {e!r}"[:200]}
    return run_ref(args, train_per_worker=TRAIN_PER_WORKER, test_per_worker=TEST_PER_WORKER, batch=BATCH)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    assert args.warmup >= 3 or args.impl == "reference" or os.environ.get("DLS_BENCH_ALLOW_SHORT"), "warmup >= 3"
    result = run_reference(args) if args.impl == "reference" else run_ours(args)
    exit_code = int(result.pop("exit_code", 0)) if isinstance(result, dict) else 0
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(result), flush=True)
    if exit_code:
        sys.exit(exit_code)


if __name__ == "__main__":
    main()
