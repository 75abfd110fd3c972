"""The remaining BASELINE.json workloads on a GPU, tiny: DenseNet-40 FedOBD with several workers
packed on one device (config 4), Transformer FedPAQ / sign_SGD on IMDB-shape tokens (config 5),
and the end-to-end data path (pinned host shards copied every epoch)."""
import pytest

pytestmark = pytest.mark.gpu


def _run(conf, **overrides):
    from distributed_learning_simulator_b200 import load_config_from_file, train
    from distributed_learning_simulator_b200.training import get_last_training_result

    config = load_config_from_file(conf, overrides=overrides)
    config.round_timeout = 300
    train(config)
    return get_last_training_result()


def test_fed_obd_densenet40_packed_workers():
    from distributed_learning_simulator_b200.engine import graph_step
    from distributed_learning_simulator_b200.ops import bn, reduce as R

    before = (graph_step.REPLAY_COUNTER[0], bn.STATS["fused_fwd"], R.STATS["packed"] + R.STATS["f32"])
    res = _run("large_scale/fed_obd/cifar10.yaml", worker_number=4, round=2, epoch=1, batch_size=32, use_amp=True,
               dataset_kwargs={"train_size": 512, "test_size": 128},
               algorithm_kwargs={"second_phase_epoch": 1, "dropout_rate": 0.3, "random_client_number": 3})
    assert sorted(res["performance"]) == [1, 2, 3]
    assert graph_step.REPLAY_COUNTER[0] > before[0], "local SGD must run as CUDA-graph replays"
    assert bn.STATS["fused_fwd"] > before[1], "DenseNet BN+ReLU must use the fused kernel"


@pytest.mark.parametrize("conf", ["fed_paq/imdb.yaml", "sign_sgd/imdb.yaml"])
def test_transformer_imdb_shape(conf):
    extra = {"algorithm_kwargs": {"random_client_number": 2}} if "paq" in conf else {}
    res = _run(conf, worker_number=2, round=1, epoch=1, batch_size=16,
               dataset_kwargs={"train_size": 96, "test_size": 64, "max_len": 300, "vocab": 2000},
               model_kwargs={"max_len": 300, "word_vector_name": "glove.6B.100d", "num_encoder_layer": 2,
                             "d_model": 100, "nhead": 5}, **extra)
    assert 1 in res["performance"]


def test_pinned_host_storage_counts_h2d_bytes():
    from distributed_learning_simulator_b200.engine import graph_step

    before = graph_step.H2D_BYTES[0]
    res = _run("fed_avg/mnist.yaml", worker_number=2, round=2, epoch=2, batch_size=32,
               dataset_kwargs={"train_size": 640, "test_size": 128}, dataloader_kwargs={"storage": "pinned_host"})
    assert sorted(res["performance"]) == [1, 2]
    per_worker_shard = 320 * (28 * 28 * 4 + 8)
    assert graph_step.H2D_BYTES[0] - before == 2 * 2 * 2 * per_worker_shard  # workers x rounds x epochs
