"""Failure detection: the reference's server polls forever when a worker dies
(server/server.py:64-85 has no timeout).  Here a role that raises aborts the topology so every peer
wakes up with an error, and a silent peer trips the round watchdog (``round_timeout``)."""
import time

import pytest

from distributed_learning_simulator_b200 import load_config_from_file, train


def _config(small_overrides, **extra):
    config = load_config_from_file("fed_avg/mnist.yaml", overrides={
        **small_overrides, "round": 2, "epoch": 1, "worker_number": 3, "learning_rate": 0.05, **extra})
    return config


def test_a_failing_worker_fails_the_task_instead_of_hanging(small_overrides, monkeypatch):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    from distributed_learning_simulator_b200.worker.aggregation_worker import AggregationWorker

    original = AggregationWorker._get_sent_data

    def broken(self):
        if self.worker_id == 1 and self._round_num == 1:
            raise ValueError("injected fault in worker 1")
        return original(self)

    monkeypatch.setattr(AggregationWorker, "_get_sent_data", broken)
    config = _config(small_overrides)
    config.round_timeout = 60
    start = time.monotonic()
    with pytest.raises(RuntimeError) as err:
        train(config)
    assert time.monotonic() - start < 50, "peers must be woken by the abort, not by the watchdog"
    assert "injected fault" in str(err.value.__cause__)


def test_a_silent_worker_trips_the_round_watchdog(small_overrides, monkeypatch):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    from distributed_learning_simulator_b200.worker.aggregation_worker import AggregationWorker

    original = AggregationWorker._get_sent_data

    def stuck(self):
        if self.worker_id == 2:
            time.sleep(12)  # far longer than the watchdog below
        return original(self)

    monkeypatch.setattr(AggregationWorker, "_get_sent_data", stuck)
    config = _config(small_overrides)
    config.round_timeout = 3
    start = time.monotonic()
    with pytest.raises(RuntimeError) as err:
        train(config)
    assert time.monotonic() - start < 30
    assert isinstance(err.value.__cause__, (TimeoutError, RuntimeError))
