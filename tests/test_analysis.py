"""Analysis tools against a real (tiny) run: log scraping, Session, graph analyzer, ModuleDiff."""
import os


from distributed_learning_simulator_b200 import load_config_from_file, train
from distributed_learning_simulator_b200.analysis import GraphSession, ModuleDiff, Session
from distributed_learning_simulator_b200.analysis import analyze_log, analyze_round, graph_exp_analyzer
from distributed_learning_simulator_b200.training import get_last_training_result


def test_log_scraping_and_session(monkeypatch, tmp_path):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    config = load_config_from_file("fed_dropout_avg/cifar10.yaml", overrides={
        "model_name": "LeNet5", "dataset_name": "MNIST", "worker_number": 2, "round": 2, "epoch": 1,
        "dataset_kwargs": {"train_size": 200, "test_size": 80},
        "algorithm_kwargs": {"dropout_rate": 0.3, "random_client_number": 2}})
    train(config)
    res = get_last_training_result()
    from distributed_learning_simulator_b200.engine import log as log_mod

    for h in list(log_mod.get_logger().handlers):
        h.flush()
    log_file = res["save_dir"].replace(os.sep + "session" + os.sep, os.sep + "log" + os.sep) + ".log"
    assert os.path.isfile(log_file)
    lines = analyze_log._read(log_file)
    assert analyze_log.final_test_accuracy(lines, "fed_dropout_avg") is not None
    assert set(analyze_log.worker_train_accuracy(lines, 2)) == {0, 1}
    amount = analyze_log.compute_data_amount([log_file], config, parameter_number=61706)
    assert amount["msg_num"] == 2 * 2 * 2 + 2 and amount["data_amount"]["mean"] > 0
    session = Session(os.path.join(res["save_dir"], "server"))
    assert session.rounds == [1, 2] and 0 <= session.mean_test_acc <= 1
    assert set(session.worker_data) == {"worker_0", "worker_1"}
    frames = analyze_round.extract_data(os.path.join(str(tmp_path), "session"), "fed_dropout_avg", {})
    assert len(frames["test_accuracy"]) == 2
    assert any(p.endswith("test_accuracy.csv") for p in analyze_round.plot(frames, str(tmp_path)))


def test_graph_session_summary(monkeypatch, tmp_path):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    config = load_config_from_file("fed_aas/dblp.yaml", overrides={
        "round": 1, "epoch": 1, "worker_number": 2, "dataset_kwargs": {"num_nodes": 200, "features": 8, "classes": 3}})
    train(config)
    res = get_last_training_result()
    summary = graph_exp_analyzer.summarize(GraphSession(os.path.join(res["save_dir"], "server")))
    assert summary["distributed_algorithm"] == "fed_aas" and "training_node_cnt" in summary
    df = graph_exp_analyzer.write(summary, str(tmp_path))
    assert len(df) == 1 and os.path.isfile(tmp_path / "exp.txt")


def test_module_diff_reports_changed_modules():
    from distributed_learning_simulator_b200 import load_config_from_dict

    config = load_config_from_dict({"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
                                    "worker_number": 1, "round": 1, "epoch": 1,
                                    "dataset_kwargs": {"train_size": 64, "test_size": 32}})
    trainer = config.create_trainer()
    hook = ModuleDiff(delta=0.01)
    trainer.append_hook(hook, "module_diff")
    hook._before_execute(executor=trainer)
    params = {k: v.clone() for k, v in trainer.model_util.get_parameter_dict().items()}
    params["fc3.weight"] = params["fc3.weight"] + 1.0
    trainer.load_parameter_dict(params)
    assert list(hook.last_report) == ["fc3"] and hook.last_report["fc3"] > 1.0
