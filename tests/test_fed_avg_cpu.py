"""BASELINE.json config 1: fed_avg LeNet-5 on MNIST-shape data, 4 workers, 2 rounds, CPU."""
import json
import os

import torch

from distributed_learning_simulator_b200 import load_config_from_file, train
from distributed_learning_simulator_b200.training import get_last_training_result


def test_fed_avg_lenet5_cpu(small_overrides, monkeypatch):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    config = load_config_from_file("fed_avg/mnist.yaml", overrides={
        **small_overrides, "round": 2, "epoch": 1, "worker_number": 4, "learning_rate": 0.05})
    train(config)
    result = get_last_training_result()
    perf = result["performance"]
    assert sorted(perf) == [1, 2]
    assert all(0.0 <= v["test_accuracy"] <= 1.0 for v in perf.values())
    server_dir = os.path.join(config.save_dir, "server")
    with open(os.path.join(server_dir, "round_record.json"), encoding="utf8") as f:
        record = json.load(f)
    assert set(record) == {"1", "2"}
    assert os.path.isfile(os.path.join(server_dir, "config.pkl"))
    assert os.path.isfile(os.path.join(server_dir, "best_global_model.pk"))
    for w in range(4):
        assert os.path.isfile(os.path.join(config.save_dir, f"worker_{w}", "hyper_parameter.pk"))
    import pickle

    with open(os.path.join(server_dir, "best_global_model.pk"), "rb") as f:
        model = pickle.load(f)
    assert sum(v.numel() for v in model.values()) == 61706
    assert all(isinstance(v, torch.Tensor) for v in model.values())
