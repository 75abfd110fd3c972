"""Concurrent independent FL tasks through the task-id API -- the reference's only pytest
(/root/reference/simulation_lib/test/test_concurrent.py:11-46): shared practitioners, several
``train(config, practitioners)`` calls in flight, results drained with get_training_result."""
import copy

from distributed_learning_simulator_b200 import Practitioner, get_training_result, load_config_from_file, train
from distributed_learning_simulator_b200.sampler import get_dataset_collection_sampler


def test_concurrent_training(monkeypatch) -> None:
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    config = load_config_from_file("fed_avg/mnist.yaml",
                                   overrides={"dataset_kwargs": {"train_size": 300, "test_size": 100}})
    config.hyper_parameter_config.epoch = 1
    config.round = 1
    config.worker_number = 3
    practitioners = set()
    sampler = get_dataset_collection_sampler(
        name=config.dataset_sampling, dataset_collection=config.create_dataset_collection(),
        part_number=config.worker_number, **config.dataset_sampling_kwargs)
    for practitioner_id in (7, 3, 11):  # ids need not be 0..n-1: they are re-numbered per task
        practitioner = Practitioner(practitioner_id=practitioner_id)
        practitioner.set_sampler(name=config.dc_config.dataset_name, sampler=sampler)
        practitioners.add(practitioner)
    task_ids = set()
    for _ in range(3):
        task_id = train(config=config, practitioners=practitioners)
        assert task_id is not None
        task_ids.add(task_id)
    assert len(task_ids) == 3
    while task_ids:
        for task_id in copy.copy(task_ids):
            result = get_training_result(task_id=task_id, timeout=300)
            assert result is not None and 1 in result["performance"]
            task_ids.remove(task_id)


def test_random_label_iid_sampler():
    import random

    from distributed_learning_simulator_b200.engine.dataset import create_dataset_collection
    from distributed_learning_simulator_b200.engine.ml_type import MachineLearningPhase

    random.seed(0)
    dc = create_dataset_collection("MNIST", {"train_size": 2000, "test_size": 400})
    sampler = get_dataset_collection_sampler("random_label_iid", dc, part_number=6, sampled_class_number=4)
    seen = set()
    for part in range(6):
        d = dc.clone()
        sampler.set_dataset_collection(d)
        sampler.sample(part)
        labels = d.get_dataset_util(MachineLearningPhase.Training).get_labels()
        assert labels <= set(sampler.assigned_labels[part]) and len(labels) <= 4
        seen |= labels
    assert seen == set(range(10))
    total = sum(len(v) for v in sampler._dataset_indices[MachineLearningPhase.Training].values())
    assert total == 2000


def test_capture_guard_excludes_eager_sections():
    """device.CaptureGuard: shared sections overlap each other, are re-entrant, and never overlap an
    exclusive (capture) section; a waiting capture is not starved by new sharers."""
    import threading
    import time

    from distributed_learning_simulator_b200.engine.device import CaptureGuard

    guard = CaptureGuard()
    state = {"shared": 0, "max_shared": 0, "exclusive": 0, "violations": 0}
    lock = threading.Lock()

    def eager(n: int) -> None:
        for _ in range(n):
            with guard.shared():
                with guard.shared():  # nested evaluate inside an eager epoch
                    with lock:
                        state["shared"] += 1
                        state["max_shared"] = max(state["max_shared"], state["shared"])
                        state["violations"] += state["exclusive"]
                    time.sleep(0.002)
                    with lock:
                        state["shared"] -= 1

    def capture(n: int) -> None:
        for _ in range(n):
            with guard.exclusive():
                with lock:
                    state["exclusive"] += 1
                    state["violations"] += state["shared"] + (state["exclusive"] - 1)
                time.sleep(0.003)
                with lock:
                    state["exclusive"] -= 1
            time.sleep(0.001)

    threads = [threading.Thread(target=eager, args=(40,)) for _ in range(4)]
    threads += [threading.Thread(target=capture, args=(10,)) for _ in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=60)
    assert not any(t.is_alive() for t in threads)
    assert state["violations"] == 0 and state["max_shared"] >= 2


def test_capture_guard_released_breaks_wait_cycles():
    """A role that blocks on another role while inside an eager section gives its hold up: a pending
    capture (which stops NEW sharers) then cannot close a wait cycle."""
    import threading
    import time

    from distributed_learning_simulator_b200.engine.device import CaptureGuard

    guard = CaptureGuard()
    answered = threading.Event()
    order = []

    def waiting_worker() -> None:           # holds shared, then waits for the peer's answer
        with guard.shared():
            with guard.released():
                assert answered.wait(timeout=20)
            order.append("worker resumed")

    def capturer() -> None:
        time.sleep(0.1)
        with guard.exclusive():
            order.append("capture")
            time.sleep(0.05)

    def peer() -> None:                     # needs the guard shared before it can answer
        time.sleep(0.2)
        with guard.shared():
            order.append("peer evaluated")
        answered.set()

    threads = [threading.Thread(target=f) for f in (waiting_worker, capturer, peer)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=30)
    assert not any(t.is_alive() for t in threads), "wait cycle"
    assert order.index("capture") < order.index("worker resumed") and "peer evaluated" in order
