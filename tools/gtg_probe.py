#!/usr/bin/env python3
"""Per-round cost of the GTG-Shapley workload: coalition evaluations, Shapley time, test accuracy."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DLS_OUTPUT_ROOT", "/tmp/dls_probe")
from distributed_learning_simulator_b200 import load_config_from_dict, train  # noqa: E402
from distributed_learning_simulator_b200.algorithm.shapley_value import gtg_shapley_value as gtg  # noqa: E402
from distributed_learning_simulator_b200.training import get_last_training_result  # noqa: E402

rows = []
orig = gtg.GTGShapleyValue.compute


def compute(self, round_number):
    t0 = time.time()
    orig(self, round_number)
    rows.append({"round": round_number, "evaluations": self.evaluations, "seconds": round(time.time() - t0, 3),
                 "v_all": round(self.last_round_metric, 4)})


gtg.GTGShapleyValue.compute = compute
kwargs = json.loads(os.environ.get("ALGORITHM_KWARGS", "{}"))
train(load_config_from_dict({
    "dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "GTG_shapley_value",
    "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 10,
    "batch_size": 64, "epoch": 1, "learning_rate": 0.01, "log_level": "WARNING", "round": int(os.environ.get("ROUNDS", "6")),
    "algorithm_kwargs": kwargs, "dataset_kwargs": {"train_size": 20000, "test_size": 4000}}))
from distributed_learning_simulator_b200.ops import subset_eval  # noqa: E402

for r in rows:
    print(json.dumps(r))
print(json.dumps(subset_eval.STATS))
print(json.dumps({k: round(v, 1) for k, v in subset_eval.STAGE_MS.items()}))
print(json.dumps({k: round(v["test_accuracy"], 4) for k, v in get_last_training_result()["performance"].items()}))
