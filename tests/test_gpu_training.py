"""GPU integration: the CUDA-graph training path and the fused aggregation inside a real
federated run, checked against the eager/torch path."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _conf(**kw):
    base = {"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
            "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 2,
            "batch_size": 32, "round": 2, "epoch": 2, "learning_rate": 0.05,
            "dataset_kwargs": {"train_size": 700, "test_size": 200}}
    base.update(kw)
    return base


def test_graph_step_matches_eager():
    """Same data order, same init: captured step == eager torch.optim.SGD step."""
    from distributed_learning_simulator_b200 import load_config_from_dict
    from distributed_learning_simulator_b200.engine.ml_type import MachineLearningPhase

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    finals = []
    for use_graph in (True, False):
        config = load_config_from_dict(_conf(use_cuda_graph=use_graph))
        trainer = config.create_trainer()
        trainer.dataset_collection.remove_dataset(MachineLearningPhase.Validation)
        trainer.set_seed(123)
        trainer.set_device(torch.device("cuda", 0))
        trainer.train()
        trainer.wait_stream()
        finals.append(trainer.model_util.get_parameter_list().clone())
        m = trainer.performance_metric.get_epoch_metrics(2)
        assert 0 <= m["accuracy"] <= 1
        if use_graph:
            assert trainer._graph_step is not None and trainer._graph_step.replays == 2 * 22
    torch.testing.assert_close(finals[0], finals[1], rtol=2e-4, atol=2e-5)


def test_fed_avg_on_gpu_uses_fused_paths():
    from distributed_learning_simulator_b200 import load_config_from_dict, train
    from distributed_learning_simulator_b200.ops import runtime
    from distributed_learning_simulator_b200.training import get_last_training_result

    before = runtime.launch_count()
    config = load_config_from_dict(_conf(worker_number=3, learning_rate=0.1))
    train(config)
    perf = get_last_training_result()["performance"]
    assert sorted(perf) == [1, 2]
    assert perf[2]["test_accuracy"] > 0.15
    assert runtime.launch_count() > before


def test_smoke_entry():
    import __graft_entry__ as entry

    entry.smoke()
