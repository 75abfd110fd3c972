"""Pure-CPU unit tests of the pieces the reference never tests: closed-form FedAvg
with full and delta uploads, quantizer round-trip bounds / unbiasedness / bit accounting, samplers."""
import math

import torch

from distributed_learning_simulator_b200.engine.model_util import ParameterLayout


def _layout(numels):
    return ParameterLayout([f"t{i}" for i in range(len(numels))], [torch.Size([n]) for n in numels])


def test_fed_avg_closed_form_full_and_delta_uploads():
    """old + sum_i n_i (w_i - old) / sum n_i, whether workers send parameters or deltas, fp64 accumulation
    (reference: algorithm/fed_avg_algorithm.py:39-52, message.py:41-49)."""
    from distributed_learning_simulator_b200.algorithm.fed_avg_algorithm import FedAVGAlgorithm
    from distributed_learning_simulator_b200.message import DeltaParameterMessage, ParameterMessage

    g = torch.Generator().manual_seed(0)
    old = {"a.weight": torch.randn(4, 3, generator=g), "a.bias": torch.randn(4, generator=g)}
    sizes = [10, 30, 60]
    new = [{k: v + torch.randn(v.shape, generator=g) for k, v in old.items()} for _ in sizes]
    want = {k: sum(n * w[k].double() for n, w in zip(sizes, new)) / sum(sizes) for k in old}
    for mode in ("full", "delta", "mixed"):
        algo = FedAVGAlgorithm()
        for i, (n, w) in enumerate(zip(sizes, new)):
            as_delta = mode == "delta" or (mode == "mixed" and i % 2 == 1)
            if as_delta:
                msg = DeltaParameterMessage(dataset_size=n, delta_parameter={k: w[k] - old[k] for k in old})
            else:
                msg = ParameterMessage(dataset_size=n, parameter={k: v.clone() for k, v in w.items()})
            algo.process_worker_data(i, msg, save_dir="", old_parameter_dict=old)
        result = algo.aggregate_worker_data()
        assert isinstance(result, ParameterMessage)
        for k in old:
            torch.testing.assert_close(result.parameter[k].double(), want[k], rtol=1e-6, atol=1e-6)
        assert not result.end_training


def test_fed_avg_skips_absent_workers_and_propagates_end_training():
    from distributed_learning_simulator_b200.algorithm.fed_avg_algorithm import FedAVGAlgorithm
    from distributed_learning_simulator_b200.message import ParameterMessage

    old = {"w": torch.zeros(5)}
    algo = FedAVGAlgorithm()
    algo.process_worker_data(0, ParameterMessage(dataset_size=1, parameter={"w": torch.ones(5)}, end_training=True),
                             save_dir="", old_parameter_dict=old)
    algo.process_worker_data(1, None, save_dir="", old_parameter_dict=old)  # not selected this round
    algo.process_worker_data(2, ParameterMessage(dataset_size=3, parameter={"w": torch.full((5,), 5.0)}),
                             save_dir="", old_parameter_dict=old)
    result = algo.aggregate_worker_data()
    torch.testing.assert_close(result.parameter["w"], torch.full((5,), 4.0))
    assert result.end_training


def test_stochastic_quantization_is_bounded_and_unbiased():
    """8-bit stochastic quantization per tensor (fed_paq / fed_obd_sq): |x - deq(q(x))| <= max|x| / 255 and
    E[deq(q(x))] = x; zero tensors and signs survive."""
    from distributed_learning_simulator_b200.algorithm.quantization.stochastic import sq_dequantize_flat, sq_quantize_flat

    layout = _layout([1000, 7, 64, 1])
    g = torch.Generator().manual_seed(1)
    flat = torch.randn(layout.total, generator=g) * torch.tensor([3.0] * 1000 + [0.0] * 7 + [0.01] * 64 + [5.0])
    mean = torch.zeros_like(flat)
    reps = 200
    for r in range(reps):
        q = sq_quantize_flat(flat, layout, 255, generator=torch.Generator().manual_seed(100 + r))
        deq = sq_dequantize_flat(q)
        for i in range(len(layout.names)):
            seg = slice(layout.offsets[i], layout.offsets[i + 1])
            bound = float(flat[seg].abs().max()) / 255 + 1e-7
            assert float((deq[seg] - flat[seg]).abs().max()) <= bound
        assert torch.equal(torch.sign(deq) * (deq != 0), torch.sign(flat) * (deq != 0))
        mean += deq / reps
    assert q.payload.dtype == torch.uint8 and q.payload.numel() == layout.total
    step = 3.0 * 4 / 255  # loose per-element step of the big tensor
    assert float((mean - flat).abs().max()) < 4 * step / math.sqrt(reps) + 1e-3
    assert float(mean[1000:1007].abs().max()) == 0.0


def test_nnadq_round_trip_error_follows_the_chosen_bit_width():
    """NNADQ (fed_obd): per tensor, error <= step / 2 with step = range / (2^bits - 1); a larger weight
    (tolerance) buys fewer bits and a smaller payload."""
    from distributed_learning_simulator_b200.algorithm.quantization.deterministic import (nnadq_dequantize_flat,
                                                                                          nnadq_quantize_flat)

    layout = _layout([4096, 33, 5, 1024])
    g = torch.Generator().manual_seed(2)
    flat = torch.randn(layout.total, generator=g)
    flat[4096 + 33: 4096 + 38] = 0.25  # a constant tensor: range 0
    sizes = {}
    for weight in (0.001, 0.01, 0.1):
        q = nnadq_quantize_flat(flat, layout, weight)
        deq = nnadq_dequantize_flat(q)
        for i in range(len(layout.names)):
            seg = slice(layout.offsets[i], layout.offsets[i + 1])
            rng = float(flat[seg].max() - flat[seg].min())
            bits = int(q.bits[i])
            assert 1 <= bits <= 16
            step = rng / (2 ** bits - 1) if rng > 0 else 0.0
            assert float((deq[seg] - flat[seg]).abs().max()) <= step / 2 + 1e-6, (weight, i, bits)
        sizes[weight] = int(q.payload.numel())
        assert int(q.byte_offsets[-1]) == q.payload.numel()
    assert sizes[0.001] > sizes[0.01] > sizes[0.1]
    assert sizes[0.1] < layout.total * 4 / 3  # well under fp32


def test_random_label_iid_split_covers_every_label_and_partitions_the_samples():
    """``random_label_iid`` (reference: sampler/base.py:9-46): each part holds exactly the requested number of
    classes, every class is held by somebody, and no training sample is given to two parts."""
    import random

    from distributed_learning_simulator_b200 import load_config_from_dict
    from distributed_learning_simulator_b200.engine.ml_type import MachineLearningPhase

    random.seed(3)
    config = load_config_from_dict({
        "dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg", "worker_number": 5,
        "round": 1, "epoch": 1, "log_level": "WARNING", "dataset_sampling": "random_label_iid",
        "dataset_sampling_kwargs": {"sampled_class_number": 4},
        "dataset_kwargs": {"train_size": 2000, "test_size": 100}})
    practitioners = config.create_practitioners()
    seen: set = set()
    labels_seen: set = set()
    for p in sorted(practitioners, key=lambda p: p.worker_id):
        trainer = p.create_trainer(config)
        util = trainer.dataset_collection.get_dataset_util(MachineLearningPhase.Training)
        indices = set(int(i) for i in util.indices)
        assert indices and not (indices & seen)
        seen |= indices
        labels = set(int(t) for t in util.targets[util.indices])
        assert len(labels) <= 4
        labels_seen |= labels
    assert labels_seen == set(range(10))


def test_round_ledger_files_plateau_rule_and_checkpoint_round_trip(tmp_path):
    """The aggregation server's bookkeeping on its own (server/aggregation_server.py): ``round_record.json`` follows
    every record, the best model is rewritten only on improvement, the initial-model record moves to key 0, the
    early-stop rule fires after five checks without a 0.001 improvement (reference aggregation_server.py:160-184),
    and a checkpoint restores the history."""
    import json
    import pickle

    import torch

    from distributed_learning_simulator_b200.server.aggregation_server import _CheckpointStore, _RoundLedger

    ledger = _RoundLedger(str(tmp_path))
    model = {"w": torch.ones(3)}
    ledger.add(1, {"accuracy": 0.10, "loss": 2.0}, model)
    ledger.rename(1, 0)                                   # metric of the initial model (need_init_performance)
    assert json.load(open(tmp_path / "round_record.json")) == {"0": {"test_accuracy": 0.10, "test_loss": 2.0}}
    assert pickle.load(open(tmp_path / "best_global_model.pk", "rb"))["w"].tolist() == [1.0, 1.0, 1.0]
    ledger.add(1, {"accuracy": 0.50, "loss": 1.0}, {"w": torch.full((3,), 2.0)})
    assert pickle.load(open(tmp_path / "best_global_model.pk", "rb"))["w"][0] == 2.0
    assert ledger.on_plateau(1) is False                  # 0.50 is an improvement over the initial reference 0
    fired = []
    for key in range(2, 8):
        ledger.add(key, {"accuracy": 0.5005, "loss": 1.0}, {"w": torch.full((3,), float(key))})
        fired.append(ledger.on_plateau(key))
    assert fired == [False, False, False, False, True, True]   # five flat checks in a row
    assert pickle.load(open(tmp_path / "best_global_model.pk", "rb"))["w"][0] == 2.0   # 0.5005 first seen at key 2
    try:
        ledger.add(3, {"accuracy": 0.9, "loss": 0.1}, model)
        raise AssertionError("a round is recorded once")
    except KeyError:
        pass

    store = _CheckpointStore(str(tmp_path), interval=2)
    assert [r for r in range(1, 7) if store.due(r)] == [2, 4, 6] and not _CheckpointStore(str(tmp_path), 0).due(4)
    store.save(4, {"w": torch.arange(3.0)}, ledger)
    state = _CheckpointStore.load(str(tmp_path))          # a directory resolves to server/checkpoint.pk
    assert state["round"] == 4 and state["parameter"]["w"].tolist() == [0.0, 1.0, 2.0]
    restored = _RoundLedger(str(tmp_path / "elsewhere"))
    restored.restore(state)
    assert restored.records == ledger.records and restored.plateau_length == ledger.plateau_length
    assert restored.best_saved == ledger.best_saved


def test_server_listening_loop_sweep_modes_and_watchdog():
    """``Server._listen`` (server/server.py) with a scripted endpoint: in one-message-per-sweep mode a worker that has
    already been heard is not listened to again until the sweep is complete; in any-time mode (sign_SGD) everybody is
    expected on every wait; an empty wait raises instead of spinning (reference server.py:64-85 polls for ever)."""
    import pytest

    from distributed_learning_simulator_b200.server.server import Server

    class _Endpoint:
        worker_num = 3

        def __init__(self, script):
            self.script, self.asked = list(script), []

        def wait_for_data(self, expected, timeout=None):
            self.asked.append(sorted(expected))
            return self.script.pop(0) if self.script else set()

        def get(self, worker_id):
            return f"m{worker_id}"

    def listener(script, per_sweep, stop_after):
        server = Server.__new__(Server)                       # the loop needs only these attributes
        server._endpoint = _Endpoint(script)
        server._one_message_per_sweep = per_sweep
        server._Executor__config = type("C", (), {"worker_number": 3, "round_timeout": 5})()
        heard: list = []
        server._stopped = lambda: len(heard) >= stop_after
        return server, heard

    server, heard = listener([{2}, {0, 1}, {1}], per_sweep=True, stop_after=4)
    for worker_id, message in server._listen():
        heard.append((worker_id, message))
    assert heard == [(2, "m2"), (0, "m0"), (1, "m1"), (1, "m1")]
    assert server._endpoint.asked == [[0, 1, 2], [0, 1], [0, 1, 2]]   # worker 2 is not awaited twice in a sweep

    server, heard = listener([{2}, {2}, {0}], per_sweep=False, stop_after=3)
    for worker_id, message in server._listen():
        heard.append((worker_id, message))
    assert [w for w, _ in heard] == [2, 2, 0] and server._endpoint.asked == [[0, 1, 2]] * 3

    server, heard = listener([{0}], per_sweep=True, stop_after=9)
    with pytest.raises(TimeoutError):
        for worker_id, message in server._listen():
            heard.append((worker_id, message))
    assert heard == [(0, "m0")]
