#!/usr/bin/env python3
"""fed_obd DenseNet-40 with many workers per GPU (debugging aid for BASELINE config 4)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DLS_OUTPUT_ROOT", "/tmp/dls_obd")
from distributed_learning_simulator_b200 import load_config_from_dict, train  # noqa: E402

if __name__ == "__main__":
    w = int(os.environ.get("WORKERS", "12"))
    train(load_config_from_dict({
        "dataset_name": "CIFAR10", "model_name": os.environ.get("MODEL", "densenet40"),
        "distributed_algorithm": os.environ.get("ALGO", "fed_obd"),
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": w,
        "batch_size": 64, "round": int(os.environ.get("ROUNDS", "3")), "epoch": 1, "learning_rate": 0.1, "use_amp": True,
        "log_level": "WARNING",
        "endpoint_kwargs": {"server": {"weight": 0.001}, "worker": {"weight": 0.001}},
        "algorithm_kwargs": {"second_phase_epoch": 1, "dropout_rate": 0.3, "random_client_number": w},
        "dataset_kwargs": {"train_size": 256 * w, "test_size": 1000},
    }))
    import torch

    if torch.cuda.is_available():
        free, total = torch.cuda.mem_get_info()
        print(f"workers {w}: device memory in use {(total - free) / 2**30:.1f} GiB of {total / 2**30:.0f} GiB, "
              f"torch reserved {torch.cuda.memory_reserved() / 2**30:.1f} GiB, "
              f"peak allocated {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
    print("done")
