#!/usr/bin/env python3
"""Per-phase device timing (DLS_TRACE=1) of FedOBD DenseNet-40 (8 workers packed on one GPU; default) or of
bench.py's flagship config (CONFIG=flagship)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DLS_OUTPUT_ROOT", "/tmp/dls_probe")
os.environ["DLS_TRACE"] = "1"
from distributed_learning_simulator_b200 import load_config_from_dict, train  # noqa: E402
from distributed_learning_simulator_b200.training import get_last_training_result  # noqa: E402
from distributed_learning_simulator_b200.utils import tracing  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_configs import CONFIGS  # noqa: E402

if os.environ.get("CONFIG", "fed_obd") == "flagship":  # bench.py's config, one worker
    conf = {"dataset_name": "CIFAR10", "model_name": "resnet18", "distributed_algorithm": "fed_avg",
            "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 1,
            "batch_size": 64, "epoch": 1, "learning_rate": 0.1, "use_amp": True, "log_level": "WARNING",
            "dataset_kwargs": {"train_size": 6250, "test_size": 1250}}
else:
    conf = dict(CONFIGS["fed_obd_densenet40_8w"])
conf["round"] = int(os.environ.get("ROUNDS", "4"))
train(load_config_from_dict(conf))
res = get_last_training_result()
trace_path = os.path.join(res["save_dir"], "server", "trace.json")
print(open(trace_path).read() if os.path.exists(trace_path) else json.dumps(tracing.summary()))
