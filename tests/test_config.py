"""Hydra-compatible CLI grammar, global.yaml merge, session naming, tracing."""
import json
import os

from distributed_learning_simulator_b200.config import (CONF_DIR, global_config, load_config, load_config_from_file,
                                                        parse_overrides)


def test_cli_overrides_with_package_prefix():
    cfg = load_config(["--config-name", "fed_avg/mnist.yaml", "++fed_avg.round=1", "++fed_avg.epoch=3",
                       "++fed_avg.worker_number=2", "++fed_avg.debug=True",
                       "++fed_avg.algorithm_kwargs.random_client_number=1"])
    assert cfg is global_config
    assert (cfg.round, cfg.hyper_parameter_config.epoch, cfg.worker_number, cfg.debug) == (1, 3, 2, True)
    assert cfg.algorithm_kwargs == {"random_client_number": 1}
    assert cfg.dc_config.dataset_name == "MNIST" and cfg.model_config.model_name == "LeNet5"
    # conf/global.yaml is merged underneath
    assert cfg.cache_transforms == "cpu" and cfg.trainer_config.hook_config.use_amp is False
    parts = cfg.save_dir.split(os.sep)
    i = parts.index("session")
    assert parts[i + 1: i + 4] == ["fed_avg", "MNIST_iid", "LeNet5"]
    assert cfg.log_file.endswith(".log") and os.sep + "log" + os.sep in cfg.log_file


def test_nested_package_and_plain_overrides():
    cfg = load_config(["--config-name", "large_scale/fed_obd/cifar10.yaml", "++large_scale.fed_obd.round=2",
                       "worker_number=7", "+endpoint_kwargs.server.weight=0.5"])
    assert cfg.round == 2 and cfg.worker_number == 7
    assert cfg.endpoint_kwargs["server"]["weight"] == 0.5 and cfg.endpoint_kwargs["worker"]["weight"] == 0.001
    assert cfg.algorithm_kwargs["random_client_number"] == 50
    assert parse_overrides(["++a.b.c=0.9", "x=[1,2]", "y=abc"], ["a"]) == {"b": {"c": 0.9}, "x": [1, 2], "y": "abc"}


def test_every_config_file_loads_and_names_a_registered_algorithm():
    import glob

    import distributed_learning_simulator_b200.method  # noqa: F401
    from distributed_learning_simulator_b200 import CentralizedAlgorithmFactory

    files = sorted(glob.glob(os.path.join(CONF_DIR, "**", "*.yaml"), recursive=True))
    assert len(files) == 54
    for path in files:
        if path.endswith("global.yaml"):
            continue
        cfg = load_config_from_file(path)
        assert CentralizedAlgorithmFactory.has_algorithm(cfg.distributed_algorithm), (path, cfg.distributed_algorithm)
        assert cfg.worker_number > 0 and cfg.round > 0
    assert sorted(CentralizedAlgorithmFactory.get_algorithm_names()) == sorted([
        "fed_avg", "fed_paq", "fed_dropout_avg", "fed_obd", "fed_obd_sq", "fed_gnn", "fed_gcn", "fed_aas",
        "multiround_shapley_value", "GTG_shapley_value", "Hierarchical_shapley_value", "sign_SGD", "single_model_afd"])


def test_tracing_writes_phase_summary(monkeypatch):
    from distributed_learning_simulator_b200 import train
    from distributed_learning_simulator_b200.training import get_last_training_result

    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    monkeypatch.setenv("DLS_TRACE", "1")
    cfg = load_config_from_file("fed_avg/mnist.yaml", overrides={
        "round": 2, "epoch": 1, "worker_number": 2, "dataset_kwargs": {"train_size": 200, "test_size": 80}})
    train(cfg)
    with open(os.path.join(get_last_training_result()["save_dir"], "server", "trace.json"), encoding="utf8") as f:
        trace = json.load(f)
    assert trace["aggregate"]["count"] == 2 and trace["evaluate"]["count"] == 2
    assert trace["local_train"]["count"] == 4 and trace["upload"]["count"] == 4
