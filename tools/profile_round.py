#!/usr/bin/env python3
"""Short federated run used under ncu: ResNet-18 fed_avg, 2 workers on one GPU, tiny shards,
so that CUDA-graph replays, the fused SGD step and the fused aggregation all appear within
a few thousand launches."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DLS_OUTPUT_ROOT", "/tmp/dls_profile")
from distributed_learning_simulator_b200 import load_config_from_dict, train  # noqa: E402

if __name__ == "__main__":
    train(load_config_from_dict({
        "dataset_name": "CIFAR10", "model_name": os.environ.get("MODEL", "resnet18"), "distributed_algorithm": "fed_avg",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR",
        "worker_number": int(os.environ.get("WORKERS", "2")), "batch_size": 64, "round": 2, "epoch": 1,
        "learning_rate": 0.1, "use_amp": True, "log_level": "WARNING",
        "dataset_kwargs": {"train_size": int(os.environ.get("TRAIN", "1024")), "test_size": 256},
    }))
    print("done")
