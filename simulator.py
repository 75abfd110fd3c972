#!/usr/bin/env python3
"""CLI entry point (same invocation as /root/reference/simulator.py:1-13):

    python3 simulator.py --config-name fed_avg/mnist.yaml ++fed_avg.round=1 ++fed_avg.worker_number=2
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from distributed_learning_simulator_b200.config import global_config, load_config  # noqa: E402
from distributed_learning_simulator_b200.training import train  # noqa: E402

if __name__ == "__main__":
    load_config()
    global_config.apply_global_config()
    train(config=global_config)
