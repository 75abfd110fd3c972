"""Every registered method runs end to end on CPU on tiny synthetic data (the semantics of
the reference's smoke scripts test.sh / other_method_test.sh, incl. partial participation
which the reference cannot run: SURVEY.md defect D1)."""
import os

import pytest
import torch

from distributed_learning_simulator_b200 import load_config_from_file, train
from distributed_learning_simulator_b200.training import get_last_training_result

TINY = {"dataset_kwargs": {"train_size": 320, "test_size": 120}, "round": 2, "epoch": 1, "batch_size": 32}


def _run(conf_file, monkeypatch, **overrides):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    merged = {**TINY, **overrides}
    if "dataset_kwargs" in overrides:
        merged["dataset_kwargs"] = {**TINY["dataset_kwargs"], **overrides["dataset_kwargs"]}
    config = load_config_from_file(conf_file, overrides=merged)
    config.round_timeout = 120
    train(config)
    res = get_last_training_result()
    return config, res


def test_fed_paq_partial_participation(monkeypatch):
    _, res = _run("fed_paq/cifar10.yaml", monkeypatch, model_name="LeNet5", dataset_name="MNIST", worker_number=3,
                  algorithm_kwargs={"random_client_number": 2})
    assert sorted(res["performance"]) == [1, 2]


def test_fed_dropout_avg(monkeypatch):
    _, res = _run("fed_dropout_avg/cifar10.yaml", monkeypatch, model_name="LeNet5", dataset_name="MNIST",
                  worker_number=2, algorithm_kwargs={"dropout_rate": 0.3, "random_client_number": 2})
    assert sorted(res["performance"]) == [1, 2]


@pytest.mark.parametrize("conf", ["fed_obd/cifar10.yaml", "fed_obd_sq/cifar100.yaml"])
def test_fed_obd_two_phases(monkeypatch, conf):
    config, res = _run(conf, monkeypatch, model_name="LeNet5", dataset_name="MNIST", worker_number=3,
                       algorithm_kwargs={"dropout_rate": 0.5, "random_client_number": 2, "second_phase_epoch": 2})
    # 2 stage-one rounds + one evaluation per stage-two epoch
    assert sorted(res["performance"]) == [1, 2, 3, 4]
    assert os.path.isfile(os.path.join(res["save_dir"], "server", "round_record.json"))


def test_obd_block_selection():
    from distributed_learning_simulator_b200.engine.model_util import ModelUtil
    from distributed_learning_simulator_b200.method.fed_obd import OpportunisticBlockDropoutAlgorithm
    from distributed_learning_simulator_b200.models.vision import DenseNet40

    model = DenseNet40(3, 10)
    util = ModelUtil(model)
    algo = OpportunisticBlockDropoutAlgorithm(dropout_rate=0.9, worker_id=1)
    algo.find_blocks(util)
    names = [n for b in algo.block_names for n in b]
    assert sorted(names) == sorted(util.get_parameter_dict())  # the blocks partition the parameters
    assert sum(1 for b in algo.block_names if len(b) == 3) == 38  # (BN, ReLU, Conv) runs
    prev = {k: v.clone() for k, v in util.get_parameter_dict().items()}
    cur = {k: v.clone() for k, v in prev.items()}
    big = algo.block_names[5]
    for n in big:
        cur[n] = cur[n] + 1.0
    scores = algo.block_scores(cur, prev)
    assert max(range(len(scores)), key=lambda i: scores[i]) == 5
    chosen = algo.select_blocks(scores)
    assert chosen[0] == 5
    total = sum(p.numel() for p in prev.values())
    assert sum(sum(cur[n].numel() for n in algo.block_names[b]) for b in chosen) <= 0.1 * total + 1


def test_fed_dropout_avg_algorithm_matches_closed_form():
    from distributed_learning_simulator_b200.message import ParameterMessage
    from distributed_learning_simulator_b200.method.fed_dropout_avg.algorithm import FedDropoutAvgAlgorithm

    algo = FedDropoutAvgAlgorithm()
    a = torch.tensor([1.0, 0.0, 3.0, 0.0])
    b = torch.tensor([3.0, 2.0, 0.0, 0.0])
    algo.process_worker_data(0, ParameterMessage(dataset_size=1, parameter={"w": a}), None, "")
    algo.process_worker_data(1, ParameterMessage(dataset_size=3, parameter={"w": b}), None, "")
    out = algo.aggregate_worker_data().parameter["w"]
    assert torch.allclose(out, torch.tensor([2.5, 2.0, 3.0, 0.0]))


def test_sign_sgd(monkeypatch):
    config, res = _run("sign_sgd/cifar10.yaml", monkeypatch, model_name="LeNet5", dataset_name="MNIST",
                       worker_number=3, round=1, epoch=2, learning_rate=0.01)
    assert sorted(res["performance"]) == [1, 2]
    assert os.path.isfile(os.path.join(res["save_dir"], "epoch_stat.json"))


def test_single_model_afd(monkeypatch):
    _, res = _run("smafd/cifar10.yaml", monkeypatch, model_name="LeNet5", dataset_name="MNIST", worker_number=3,
                  algorithm_kwargs={"random_client_number": 2, "dropout_rate": 0.3})
    assert sorted(res["performance"]) == [1, 2]


@pytest.mark.parametrize("conf,workers", [("gtg_sv/mnist.yaml", 3), ("multiround_sv/cifar10.yaml", 3),
                                          ("hierarchical_sv/mnist.yaml", 4)])
def test_shapley_methods(monkeypatch, conf, workers):
    import json

    overrides = {"model_name": "LeNet5", "dataset_name": "MNIST", "worker_number": workers}
    if "hierarchical" in conf:
        overrides["algorithm_kwargs"] = {"part_number": 2, "vp_size": 2}
    _, res = _run(conf, monkeypatch, **overrides)
    assert sorted(res["performance"]) == [0, 1, 2]
    assert sorted(res["sv"]) == [1, 2]
    assert sorted(res["sv"][1]) == list(range(workers))
    with open(os.path.join(res["save_dir"], "shapley_values.json"), encoding="utf8") as f:
        assert sorted(json.load(f)) == ["1", "2"]


def test_sign_sgd_workers_out_of_phase(monkeypatch):
    """Shards with different batch counts: one worker keeps stepping while the others already wait for the epoch
    acknowledgement, i.e. it sends two messages where they send one.  The server must keep listening to it."""
    config, res = _run("sign_sgd/cifar10.yaml", monkeypatch, model_name="LeNet5", dataset_name="MNIST", worker_number=3,
                       round=1, epoch=3, learning_rate=0.002, batch_size=64,
                       dataset_kwargs={"train_size": 1920, "test_size": 300})
    assert sorted(res["performance"]) == [1, 2, 3]
