#!/usr/bin/env python3
"""Rounds/s of the other BASELINE.json target workloads (GTG-Shapley LeNet-5, FedOBD DenseNet-40)
for this framework and for the unmodified reference (baseline/_ref on baseline/shims), one GPU.

Method (identical for both arms): run the whole program in a fresh process for R1 rounds and for
R2 rounds and report (wall(R2) - wall(R1)) / (R2 - R1) seconds per round, which cancels start-up,
data generation, graph capture and (for FedOBD) the constant second phase.  Full participation,
because the reference crashes on partial participation (SURVEY.md defects D1/D3)."""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CONFIGS = {
    "gtg_sv_lenet5_10w": {
        "dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "GTG_shapley_value",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 10,
        "batch_size": 64, "epoch": 1, "learning_rate": 0.01, "log_level": "WARNING",
        "dataset_kwargs": {"train_size": 20000, "test_size": 4000}},
    "fed_obd_densenet40_8w": {
        "dataset_name": "CIFAR10", "model_name": "densenet40", "distributed_algorithm": "fed_obd",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 8,
        "batch_size": 64, "epoch": 1, "learning_rate": 0.1, "use_amp": True, "log_level": "WARNING",
        "endpoint_kwargs": {"server": {"weight": 0.001}, "worker": {"weight": 0.001}},
        "algorithm_kwargs": {"second_phase_epoch": 1, "dropout_rate": 0.3, "random_client_number": 8},
        "dataset_kwargs": {"train_size": 8000, "test_size": 1000}},
}

OURS = r"""
import sys, json, time
sys.path.insert(0, {root!r})
import torch
from distributed_learning_simulator_b200 import load_config_from_dict, train
from distributed_learning_simulator_b200.utils.observer import add_round_observer
stamps = []
def stamp(server, round_number):
    torch.cuda.synchronize()
    stamps.append(time.perf_counter())
add_round_observer(stamp)
train(load_config_from_dict({conf!r}))
print("ROUND_STAMPS " + json.dumps(stamps))
"""

REFERENCE = r"""
import sys
sys.path.insert(0, {shims!r}); sys.path.insert(0, {ref!r})
from distributed_learning_simulator.config import load_config_from_file
from distributed_learning_simulator.training import train
if __name__ == "__main__":
    config = load_config_from_file({conf_path!r})
    config.apply_global_config()
    train(config=config)
"""


def run(arm: str, conf: dict, rounds: int, work: str) -> float:
    conf = dict(conf, round=rounds)
    extra = os.environ.get("ALGORITHM_KWARGS")  # JSON merged into algorithm_kwargs (A/B runs of our arm)
    if extra and arm == "ours":
        conf["algorithm_kwargs"] = dict(conf.get("algorithm_kwargs", {}), **json.loads(extra))
    script = os.path.join(work, f"{arm}_{rounds}.py")
    env = dict(os.environ, DLS_OUTPUT_ROOT=work, REF_LOG_LEVEL="WARNING", REF_GPUS="1")
    if arm == "ours":
        open(script, "w").write(OURS.format(root=ROOT, conf=conf))
    else:
        import yaml

        conf_path = os.path.join(work, f"ref_{rounds}.yaml")
        yaml.safe_dump(conf, open(conf_path, "w"))
        open(script, "w").write(REFERENCE.format(shims=os.path.join(ROOT, "baseline", "shims"),
                                                 ref=os.path.join(ROOT, "baseline", "_ref"), conf_path=conf_path))
    start = time.time()
    res = subprocess.run([sys.executable, script], cwd=work, env=env, capture_output=True, text=True,
                         timeout=int(os.environ.get("ARM_TIMEOUT", "900")))
    if res.returncode != 0:
        raise RuntimeError((res.stderr or res.stdout)[-600:])
    for line in res.stdout.splitlines():
        if line.startswith("ROUND_STAMPS "):
            ROUND_STAMPS[(arm, rounds)] = json.loads(line.split(" ", 1)[1])
    return time.time() - start


ROUND_STAMPS: dict = {}


def main() -> None:
    r1, r2 = int(os.environ.get("R1", "1")), int(os.environ.get("R2", "3"))
    which = os.environ.get("CONFIGS", ",".join(CONFIGS)).split(",")
    arms = os.environ.get("ARMS", "ours,reference").split(",")
    out = {}
    for name in which:
        out[name] = {}
        for arm in arms:
            work = tempfile.mkdtemp(prefix=f"bench_{name}_{arm}_")
            try:
                t1, t2 = run(arm, CONFIGS[name], r1, work), run(arm, CONFIGS[name], r2, work)
                out[name][arm] = {"wall_s": {r1: round(t1, 2), r2: round(t2, 2)},
                                  "s_per_round": round((t2 - t1) / (r2 - r1), 3),
                                  "rounds_per_s": round((r2 - r1) / max(t2 - t1, 1e-9), 4)}
                stamps = ROUND_STAMPS.get((arm, r2))
                if stamps and len(stamps) >= 2:
                    # our arm also reports the server's own round-end clock (after a device synchronize):
                    # rounds 2..R2 of the long run, free of start-up noise
                    deltas = [b - a for a, b in zip(stamps, stamps[1:])]
                    out[name][arm]["round_seconds"] = [round(d, 4) for d in deltas]
                    out[name][arm]["s_per_round_in_process"] = round(sum(deltas) / len(deltas), 4)
            except Exception as e:  # noqa: BLE001
                out[name][arm] = {"error": str(e)[-400:]}
            print(name, arm, json.dumps(out[name][arm]), flush=True)
    path = os.environ.get("OUT", "gpurun_out/bench_configs.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    json.dump(out, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
