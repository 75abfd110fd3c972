"""Runs the UNMODIFIED reference (``baseline/_ref``, installed with pip --target) on the
benchmark configuration and measures the same metric as ``bench.py``'s own arm.

The reference is driven through its own public API and stock code path:
``load_config_from_file`` -> ``global_config.apply_global_config()`` -> ``train(config)``
(/root/reference/simulator.py:11-13, simulation_lib/config.py:98-104).  It spawns its own
process pool (one process per GPU + a server process), so under ``torch.distributed.run`` only
rank 0 launches it -- over all ``--gpus`` devices -- and the other ranks just wait.

Timing: the reference offers no hook around a round, and its roles live in separate
processes, so rounds are bracketed by the completion time of the server's per-round test
evaluation (recorded, after a device synchronize, by the stand-in ``Inferencer`` into
REF_TIMING_FILE).  Round r ends when evaluation r has finished; the timed region spans rounds
warmup+1 .. warmup+steps.  Because the reference is a multi-process program this is host
wall-clock time on the server process, not CUDA-event time.
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import os, resource, sys
soft, hard = resource.getrlimit(resource.RLIMIT_NOFILE)
if soft <= 1024:
    resource.setrlimit(resource.RLIMIT_NOFILE, (min(max(hard, 4096), 65536) if hard > 0 else 65536, hard))
sys.path.insert(0, {shims!r})
sys.path.insert(0, {ref!r})
from distributed_learning_simulator.config import global_config, load_config_from_file
from distributed_learning_simulator.training import train
if __name__ == "__main__":
    config = load_config_from_file({conf!r})
    config.apply_global_config()
    train(config=config)
"""


def run(args, conf: dict, samples_per_round: int, meta: dict) -> dict:
    """``conf``: the flat config dict (written as YAML for ``load_config_from_file``); ``samples_per_round``:
    training samples all workers consume per round; ``meta``: metric / config description for the JSON line."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    result: dict = {}
    if rank == 0:
        result = _run_rank0(args, conf, samples_per_round, meta)
    if world > 1:
        import torch.distributed as dist

        if not dist.is_initialized():
            dist.init_process_group("gloo")
        dist.barrier()
    return result


def _run_rank0(args, conf: dict, samples_per_round: int, meta: dict) -> dict:
    import yaml

    n = args.gpus
    rounds = args.warmup + args.steps
    work = tempfile.mkdtemp(prefix="ref_arm_")
    conf_path = os.path.join(work, "bench.yaml")
    with open(conf_path, "wt", encoding="utf8") as f:
        yaml.safe_dump(conf, f)
    timing = os.path.join(work, "timing.txt")
    script = os.path.join(work, "run_reference.py")
    with open(script, "wt", encoding="utf8") as f:
        f.write(CHILD.format(shims=os.path.join(ROOT, "baseline", "shims"), ref=os.path.join(ROOT, "baseline", "_ref"),
                             conf=conf_path))
    env = {k: v for k, v in os.environ.items()
           if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "LOCAL_WORLD_SIZE",
                        "GROUP_RANK", "ROLE_RANK", "TORCHELASTIC_RUN_ID", "OMP_NUM_THREADS")}
    env.update(REF_TIMING_FILE=timing, REF_GPUS=str(n), REF_LOG_LEVEL="WARNING",
               PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "baseline", "shims"),
                                           os.path.join(ROOT, "baseline", "_ref")]))
    start = time.time()
    try:
        proc = subprocess.run([sys.executable, script], cwd=work, env=env, capture_output=True, text=True,
                              timeout=int(os.environ.get("REF_TIMEOUT", "1500")))
    except subprocess.TimeoutExpired:
        return {"impl": "reference", "unavailable": "reference run exceeded REF_TIMEOUT", "exit_code": 1}
    if proc.returncode != 0:
        return {"impl": "reference", "unavailable": ("reference run failed: " + (proc.stderr or proc.stdout)[-300:])
                .replace("\n", " "), "exit_code": 1}
    with open(timing, "rt", encoding="utf8") as f:
        stamps = [float(line) for line in f if line.strip()]
    if len(stamps) < rounds:
        return {"impl": "reference", "unavailable": f"expected {rounds} round evaluations, saw {len(stamps)}",
                "exit_code": 1}
    # the last `rounds` evaluations are the per-round test passes (an initial-model evaluation may precede them)
    stamps = stamps[-rounds:]
    seconds = stamps[args.warmup + args.steps - 1] - stamps[args.warmup - 1]
    samples = samples_per_round * args.steps
    config = dict(meta["config"])
    config["parallelism"] = f"reference process pool: {n} worker process(es) + server"
    config["timing"] = "host wall clock between per-round server evaluations (multi-process program)"
    return {
        "metric": meta["metric"], "value": samples / seconds, "unit": "samples/s", "n_gpus": n, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": seconds * 1e3 / args.steps, "higher_is_better": True,
        "scaling": meta["scaling"], "vs_baseline": None, "dtype": meta["dtype"], "data": "synthetic",
        "impl": "reference", "rounds_per_s": args.steps / seconds, "config": config,
        "e2e": {"value": samples / seconds, "unit": "samples/s",
                "note": "the reference always feeds batches from host memory; same number as value"},
        "total_wall_s": time.time() - start,
    }
