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


def test_round_counter_checkpoint_and_resume(small_overrides, monkeypatch):
    """Extension over the reference (warm start only): ``checkpoint_interval`` writes
    server/checkpoint.pk, ``resume_from`` continues with the next round number, the restored
    global model and the earlier rounds' records."""
    import pickle

    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    base = {**small_overrides, "epoch": 1, "worker_number": 2, "learning_rate": 0.05}
    first = load_config_from_file("fed_avg/mnist.yaml", overrides={
        **base, "round": 2, "algorithm_kwargs": {"checkpoint_interval": 1}})
    train(first)
    res1 = get_last_training_result()
    ck_path = os.path.join(res1["save_dir"], "server", "checkpoint.pk")
    with open(ck_path, "rb") as f:
        ck = pickle.load(f)
    assert ck["round"] == 2 and sorted(ck["stat"]) == [1, 2]
    assert sum(v.numel() for v in ck["parameter"].values()) == 61706

    second = load_config_from_file("fed_avg/mnist.yaml", overrides={
        **base, "round": 4, "algorithm_kwargs": {"resume_from": res1["save_dir"], "checkpoint_interval": 2}})
    train(second)
    res2 = get_last_training_result()
    assert sorted(res2["performance"]) == [1, 2, 3, 4]
    assert res2["performance"][1] == res1["performance"][1] and res2["performance"][2] == res1["performance"][2]
    with open(os.path.join(res2["save_dir"], "server", "checkpoint.pk"), "rb") as f:
        assert pickle.load(f)["round"] == 4
    # the resumed run really started from the check-pointed model: its first new round is no worse
    # than an untrained model would be, and workers ran exactly two more rounds
    assert res2["performance"][3]["test_loss"] < res1["performance"][1]["test_loss"]
