#!/usr/bin/env python3
"""Benchmarks of the BASELINE.json configurations; the default is the headline: federated rounds of
fed_avg ResNet-18 on synthetic CIFAR-10-shape data, one worker per B200 (config 2).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...          # the unmodified reference (baseline/_ref), same config
    python bench.py --config fed_obd|gtg_sv|fed_paq ...   # the other target workloads, both arms
    python bench.py --dtype fp32 ...              # the reference's stock precision (conf/global.yaml: use_amp false)

A "step" is one federated round: every (selected) worker trains its local epochs (batch 64, SGD with cosine
schedule, best-model-by-validation) on its shard, uploads, the server aggregates + broadcasts and evaluates
the new global model on the test split.  ``value`` is worker samples/s summed over all GPUs (training samples
consumed per second of the timed rounds); rounds/s is reported alongside.  fed_avg scales weakly (one worker
and one shard per GPU; N = 8 is exactly the 50 000 / 10 000 CIFAR-10 shape); the other configs keep the
reference's worker counts (100 / 10 / 50 workers packed on the N GPUs: strong scaling).

Two measurements per run, each with W untimed + K timed rounds bracketed by barrier + synchronize and timed
with CUDA events (max over ranks):
  * device-resident data  -> ``value``
  * end to end            -> ``e2e``: every round each shard is copied host->device from pinned memory and
    the round's metrics are read back device->host.
At N > 1 the warm-up rounds also check the fused cross-GPU reduction against an fp64 NCCL all-reduce of the
same rows (``agg_max_abs_err``; the run fails above 1e-5), and ``phases_median_ms`` carries the median
per-round device time of local_train / upload / aggregate / evaluate / broadcast (max over ranks).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 64
_TRANSFORMER = {"max_len": 300, "word_vector_name": "glove.6B.100d", "num_encoder_layer": 2, "d_model": 100, "nhead": 5}

# per config: base settings (reference YAML cited), workers as a function of the GPU count, shard sizes
CONFIGS = {
    # BASELINE.json config 2 (model by name; the reference ships no resnet18 YAML)
    "fed_avg": {
        "base": {"dataset_name": "CIFAR10", "model_name": "resnet18", "distributed_algorithm": "fed_avg",
                 "learning_rate": 0.1, "epoch": 1},
        "workers": lambda n: n, "train_per_worker": 6250, "test_per_worker": 1250, "scaling": "weak",
        "model": "resnet18 (CIFAR stem, 11.17M params)",
        "metric": "worker samples/sec (fed_avg ResNet-18, CIFAR-10 shape, federated rounds)"},
    # /root/reference/conf/large_scale/fed_obd/cifar10.yaml (100 workers; 1 local epoch per round here)
    "fed_obd": {
        "base": {"dataset_name": "CIFAR10", "model_name": "densenet40", "distributed_algorithm": "fed_obd",
                 "learning_rate": 0.1, "epoch": 1,
                 "endpoint_kwargs": {"server": {"weight": 0.001}, "worker": {"weight": 0.001}},
                 "algorithm_kwargs": {"second_phase_epoch": 1, "dropout_rate": 0.3, "random_client_number": 100}},
        "workers": lambda n: 100, "train_per_worker": 500, "test_total": 10000, "scaling": "strong",
        "model": "densenet40 (depth 40, growth 12)",
        "metric": "worker samples/sec (fed_obd DenseNet-40, CIFAR-10 shape, 100 workers, federated rounds)",
        "note": "full participation in both arms: the reference crashes when random_client_number < worker_number "
                "(SURVEY.md defect D1)"},
    # /root/reference/conf/gtg_sv/mnist.yaml, gtg_shapley_train.sh
    "gtg_sv": {
        "base": {"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "GTG_shapley_value",
                 "learning_rate": 0.01, "epoch": 1},
        "workers": lambda n: 10, "train_per_worker": 2000, "test_total": 4000, "scaling": "strong",
        "model": "LeNet5", "metric": "worker samples/sec (GTG-Shapley fed_avg LeNet-5, MNIST shape, 10 workers)"},
    # /root/reference/conf/fed_paq/imdb.yaml with BASELINE.json config 5's 50 workers
    "fed_paq": {
        "base": {"dataset_name": "imdb", "model_name": "TransformerClassificationModel", "distributed_algorithm": "fed_paq",
                 "learning_rate": 0.01, "epoch": 1, "dataset_kwargs": {"max_len": 300}, "model_kwargs": dict(_TRANSFORMER)},
        "workers": lambda n: 50, "train_per_worker": 500, "test_total": 5000, "scaling": "strong",
        "model": "TransformerClassificationModel (2 layers, d_model 100, seq 300)",
        "metric": "worker samples/sec (fed_paq Transformer, IMDB shape seq 300, 50 workers)"},
}


def config_dict(name: str, n_gpus: int, rounds: int, dtype: str, storage: str | None = None) -> dict:
    spec = CONFIGS[name]
    workers = spec["workers"](n_gpus)
    conf = json.loads(json.dumps(spec["base"]))
    test = spec.get("test_total") or spec["test_per_worker"] * workers
    conf.update({"optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": workers,
                 "batch_size": BATCH, "round": rounds, "use_amp": dtype == "bf16", "log_level": "WARNING"})
    conf.setdefault("dataset_kwargs", {}).update({"train_size": spec["train_per_worker"] * workers, "test_size": test})
    if storage is not None:
        conf["dataloader_kwargs"] = {"storage": storage}
    return conf


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


def model_macs_per_sample(conf: dict) -> float | None:
    """Multiply-accumulates of one forward pass (convolutions + linear layers), counted with hooks on a CPU copy."""
    try:
        import torch
        from torch import nn

        from distributed_learning_simulator_b200 import load_config_from_dict

        small = json.loads(json.dumps(conf))
        small["dataset_kwargs"].update({"train_size": 64, "test_size": 64})
        small["worker_number"] = 1
        small.pop("dataloader_kwargs", None)
        config = load_config_from_dict(small)
        dc = config.create_dataset_collection()
        from distributed_learning_simulator_b200.models.registry import create_model

        model = create_model(config.model_config.model_name, dc, **config.model_config.model_kwargs)
        macs = [0.0]

        def hook(m, inp, out):
            if isinstance(m, nn.Conv2d):
                macs[0] += out[0].numel() * (m.in_channels // m.groups) * m.kernel_size[0] * m.kernel_size[1]
            elif isinstance(m, nn.Linear):
                macs[0] += out[0].numel() * m.in_features

        handles = [m.register_forward_hook(hook) for m in model.modules() if isinstance(m, (nn.Conv2d, nn.Linear))]
        shape = dc.spec["shape"]
        with torch.no_grad():
            if dc.spec.get("kind", "image") == "image" or len(shape) == 3:
                model.eval()(torch.zeros(1, *shape))
            else:
                return None
        for h in handles:
            h.remove()
        return float(macs[0])
    except Exception:  # noqa: BLE001 - a diagnostic, never a reason to fail the benchmark
        return None


def run_ours(args) -> dict:
    import torch

    from distributed_learning_simulator_b200 import load_config_from_dict, train
    from distributed_learning_simulator_b200.engine import graph_step
    from distributed_learning_simulator_b200.engine.metric import PerformanceMetric
    from distributed_learning_simulator_b200.ops import runtime
    from distributed_learning_simulator_b200.training import get_last_training_result
    from distributed_learning_simulator_b200.utils.clocks import ClockSampler
    from distributed_learning_simulator_b200.utils.observer import add_round_observer, remove_round_observer

    spec = CONFIGS[args.config]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1, f"launched with WORLD_SIZE={world} but --gpus {args.gpus}"
    n_workers = spec["workers"](args.gpus)
    if world == 1 and args.gpus > 1:
        assert torch.cuda.device_count() >= args.gpus
    torch.cuda.set_device(local_rank)
    runtime.require()
    rounds = args.warmup + args.steps
    if world > 1:
        os.environ["DLS_AGG_CHECK"] = str(args.warmup)  # fused reduce vs fp64 NCCL all-reduce in the warm-up rounds
    os.environ["DLS_TRACE"] = "1"
    results = {}
    agg = {"err": 0.0, "checks": 0}
    phases: dict = {}
    for mode, storage in (("device", "device"), ("e2e", "pinned_host")):
        sampler = ClockSampler(gpu_index=local_rank) if (mode == "device" and rank == 0) else None
        timer = RoundTimer(args.warmup, args.steps, sampler)
        add_round_observer(timer)
        PerformanceMetric.d2h_bytes = 0
        counters0 = (runtime.launch_count(), graph_step.CAPTURE_LAUNCHES[0], graph_step.OWN_REPLAYED_KERNELS[0])
        conf = config_dict(args.config, args.gpus, rounds, args.dtype, storage)
        try:
            config = load_config_from_dict(conf)
            config.apply_global_config()  # log level (100 workers at INFO write tens of MB of log) and the seed
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
        if mode == "device":
            phases = _phase_table(get_last_training_result().get("save_dir"), world)
            if world > 1:
                from distributed_learning_simulator_b200.parallel import context as dist_context

                ctx = dist_context.get_context()
                t = torch.tensor([ctx.agg_max_abs_err, float(ctx.agg_checks_done)], dtype=torch.float64, device="cuda")
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
                agg = {"err": float(t[0].item()), "checks": int(t[1].item())}
        selected = spec["base"].get("algorithm_kwargs", {}).get("random_client_number", n_workers)
        samples = spec["train_per_worker"] * min(selected, n_workers) * spec["base"]["epoch"] * args.steps
        results[mode] = {
            "ms": ms, "samples_per_s": samples / (ms / 1e3), "rounds_per_s": args.steps / (ms / 1e3),
            "wall_ms": (timer.wall[1] - timer.wall[0]) * 1e3, "clocks": clocks,
            "launches": launches,
            "h2d_bytes_per_step": h2d / max(rounds, 1) if storage == "pinned_host" else 0,
            "d2h_bytes_per_step": d2h / max(rounds, 1),
        }
        graph_step.H2D_BYTES[0] = 0
        os.environ["DLS_AGG_CHECK"] = "0"
    dev, e2e = results["device"], results["e2e"]
    conf = config_dict(args.config, args.gpus, rounds, args.dtype)
    macs = model_macs_per_sample(conf) if rank == 0 else None
    flops = None
    mfu = None
    if macs:
        # training: forward + input gradient + weight gradient = 3 x 2 x MACs per sample
        flops = 6.0 * macs * dev["samples_per_s"]
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            mfu = flops / 1e12 / (float(peaks["bf16_tflops_sustained"]) * args.gpus)
        except Exception:  # noqa: BLE001
            mfu = flops / 1e12 / (1400.0 * args.gpus)
    if world > 1 and agg["checks"] > 0:
        assert agg["err"] <= 1e-5, f"fused cross-GPU reduction differs from the fp64 all-reduce by {agg['err']}"
    out = {
        "metric": spec["metric"],
        "value": dev["samples_per_s"], "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev["ms"] / args.steps, "higher_is_better": True, "scaling": spec["scaling"],
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "impl": "ours",
        "rounds_per_s": dev["rounds_per_s"],
        "config": {"name": args.config, "model": spec["model"], "algorithm": spec["base"]["distributed_algorithm"],
                   "workers": n_workers, "global_batch": BATCH * n_workers,
                   "seq_len": 300 if "imdb" in str(spec["base"]) else None,
                   "local_batch": BATCH, "local_epochs_per_round": spec["base"]["epoch"],
                   "train_samples_per_worker": spec["train_per_worker"],
                   "test_samples": conf["dataset_kwargs"]["test_size"],
                   "parallelism": f"fl-dp{args.gpus} ({n_workers} worker(s) on {args.gpus} GPU(s))",
                   "l2_policy": "inputs larger than L2: every round streams each worker's shard plus model/optimizer "
                                "state (fed_avg: 77 MB + 45 MB per worker); no explicit flush"},
        "clocks": dev["clocks"],
        "e2e": {"value": e2e["samples_per_s"], "unit": "samples/s", "ms_per_step": e2e["ms"] / args.steps,
                "h2d_bytes_per_step": e2e["h2d_bytes_per_step"], "d2h_bytes_per_step": e2e["d2h_bytes_per_step"]},
        "gpu_launches": int(dev["launches"]),
        "wall_ms_per_step": dev["wall_ms"] / args.steps,
        "model_fwd_gmacs_per_sample": None if not macs else macs / 1e9,
        "train_tflops": None if flops is None else flops / 1e12,
        "mfu_of_measured_sustained_bf16": mfu,
        "phases_median_ms": phases,
    }
    if "note" in spec:
        out["config"]["note"] = spec["note"]
    if world > 1:
        out["agg_max_abs_err"] = agg["err"]
        out["agg_checks"] = agg["checks"]
    return out


def _phase_table(save_dir: str | None, world: int) -> dict:
    """Median device milliseconds per phase from the run's trace.json, max over ranks."""
    import torch

    names = ["local_train", "upload", "aggregate", "evaluate", "broadcast"]
    vals = [0.0] * len(names)
    try:
        with open(os.path.join(save_dir, "server", "trace.json"), "rt", encoding="utf8") as f:
            trace = json.load(f)
        vals = [float(trace.get(n, {}).get("median_ms", 0.0)) for n in names]
    except Exception:  # noqa: BLE001
        pass
    if world > 1:
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        vals = [float(v) for v in t.tolist()]
    return {n: round(v, 3) for n, v in zip(names, vals)}


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
    spec = CONFIGS[args.config]
    rounds = args.warmup + args.steps
    conf = config_dict(args.config, args.gpus, rounds, args.dtype)
    n_workers = spec["workers"](args.gpus)
    samples_per_round = spec["train_per_worker"] * n_workers * spec["base"]["epoch"]
    meta = {"metric": spec["metric"], "scaling": spec["scaling"], "dtype": args.dtype,
            "config": {"name": args.config, "model": spec["model"], "algorithm": spec["base"]["distributed_algorithm"],
                       "workers": n_workers, "global_batch": BATCH * n_workers,
                       "seq_len": 300 if "imdb" in str(spec["base"]) else None, "local_batch": BATCH,
                       "local_epochs_per_round": spec["base"]["epoch"], "train_samples_per_worker": spec["train_per_worker"],
                       "test_samples": conf["dataset_kwargs"]["test_size"]}}
    if "note" in spec:
        meta["config"]["note"] = spec["note"]
    return run_ref(args, conf, samples_per_round, meta)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="fed_avg", choices=sorted(CONFIGS))
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
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
