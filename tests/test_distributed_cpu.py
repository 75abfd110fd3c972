"""One process per "GPU" with the gloo backend on CPU, world_size 2: server replicas must
agree with each other and with the single-process run (same seed => same decisions)."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import json, os, sys
sys.path.insert(0, {root!r})
from distributed_learning_simulator_b200 import load_config_from_dict, train
from distributed_learning_simulator_b200.training import get_last_training_result
conf = {conf!r}
config = load_config_from_dict(conf)
train(config)
res = get_last_training_result()
out = os.path.join({out!r}, "rank_%s.json" % os.environ.get("RANK", "0"))
with open(out, "w") as f:
    json.dump({{"performance": {{str(k): v for k, v in res["performance"].items()}}, "save_dir": res["save_dir"]}}, f)
"""


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(tmp_path, conf, world):
    out = tmp_path / f"world{world}"
    out.mkdir()
    script = tmp_path / f"run{world}.py"
    script.write_text(SCRIPT.format(root=ROOT, conf=conf, out=str(out)))
    env = dict(os.environ, DLS_FORCE_CPU="1", DLS_OUTPUT_ROOT=str(out), OMP_NUM_THREADS="2")
    if world == 1:
        cmd = [sys.executable, str(script)]
        env.pop("WORLD_SIZE", None)
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return [json.load(open(out / f"rank_{r}.json")) for r in range(world)]


def test_two_ranks_agree_with_single_process(tmp_path):
    conf = {"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
            "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 4,
            "batch_size": 32, "round": 2, "epoch": 1, "learning_rate": 0.05,
            "algorithm_kwargs": {"random_client_number": 3},
            "dataset_kwargs": {"train_size": 640, "test_size": 200}}
    two = _run(tmp_path, conf, 2)
    one = _run(tmp_path, conf, 1)
    assert two[0]["performance"] == two[1]["performance"]
    assert two[0]["save_dir"] == two[1]["save_dir"]
    assert sorted(two[0]["performance"]) == ["1", "2"]
    for k in ("1", "2"):
        a, b = two[0]["performance"][k], one[0]["performance"][k]
        assert abs(a["test_accuracy"] - b["test_accuracy"]) < 1e-4
        assert abs(a["test_loss"] - b["test_loss"]) < 1e-4
    server_dir = os.path.join(two[0]["save_dir"], "server")
    assert os.path.isfile(os.path.join(server_dir, "round_record.json"))



MULTI_SCRIPT = r"""
import json, os, sys
sys.path.insert(0, {root!r})
from distributed_learning_simulator_b200 import load_config_from_dict, train
from distributed_learning_simulator_b200.training import get_last_training_result
results = {{}}
for name, conf in {confs!r}.items():
    train(load_config_from_dict(conf))
    res = get_last_training_result()
    results[name] = {{str(k): v for k, v in res["performance"].items()}}
with open(os.path.join({out!r}, "multi_%s.json" % os.environ.get("RANK", "0")), "w") as f:
    json.dump(results, f)
"""

_TINY = {"optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "batch_size": 32, "round": 2,
         "epoch": 1, "learning_rate": 0.05, "dataset_name": "MNIST", "model_name": "LeNet5", "worker_number": 4,
         "dataset_kwargs": {"train_size": 480, "test_size": 160}}

_METHODS = {
    "fed_obd": {"endpoint_kwargs": {"server": {"weight": 0.01}, "worker": {"weight": 0.01}},
                "algorithm_kwargs": {"dropout_rate": 0.5, "random_client_number": 3, "second_phase_epoch": 1}},
    "fed_paq": {"algorithm_kwargs": {"random_client_number": 3}},
    "GTG_shapley_value": {"worker_number": 3},
    "sign_SGD": {"round": 1, "distribute_init_parameters": False, "learning_rate": 0.01},
    "fed_gnn": {"dataset_name": "Cora", "model_name": "TwoGCN",
                "dataset_kwargs": {"num_nodes": 240, "features": 16, "classes": 3},
                "algorithm_kwargs": {"share_feature": True, "batch_number": 2}},
}


def test_methods_under_two_ranks(tmp_path):
    """Every method family also runs with one process per device (one launch, run back to back):
    the replicated servers agree on every recorded metric."""
    confs = {name: {**_TINY, **extra, "distributed_algorithm": name} for name, extra in _METHODS.items()}
    script = tmp_path / "multi.py"
    script.write_text(MULTI_SCRIPT.format(root=ROOT, confs=confs, out=str(tmp_path)))
    env = dict(os.environ, DLS_FORCE_CPU="1", DLS_OUTPUT_ROOT=str(tmp_path), OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=420)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    r0, r1 = (json.load(open(tmp_path / f"multi_{r}.json")) for r in range(2))
    assert r0 == r1 and set(r0) == set(_METHODS)
    assert all(len(perf) >= 1 for perf in r0.values())


FAULT_SCRIPT = r"""
import os, sys
sys.path.insert(0, {root!r})
from distributed_learning_simulator_b200 import load_config_from_dict, train
from distributed_learning_simulator_b200.worker.aggregation_worker import AggregationWorker
original = AggregationWorker._get_sent_data
def broken(self):
    if self.worker_id == 1 and self._round_num == 1:
        raise ValueError("injected fault in worker 1")
    return original(self)
AggregationWorker._get_sent_data = broken
train(load_config_from_dict({conf!r}))
"""


def test_a_dead_worker_takes_the_whole_job_down(tmp_path):
    """A role that raises on one rank must not leave the other rank waiting in a collective for ever: the failing
    rank exits non-zero after a grace period and the launcher tears the job down (parallel/launch.py)."""
    import time

    conf = {**_TINY, "distributed_algorithm": "fed_avg", "round": 2}
    script = tmp_path / "fault.py"
    script.write_text(FAULT_SCRIPT.format(root=ROOT, conf=conf))
    env = dict(os.environ, DLS_FORCE_CPU="1", DLS_OUTPUT_ROOT=str(tmp_path), OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    start = time.monotonic()
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
    assert res.returncode != 0
    assert time.monotonic() - start < 200
    assert "injected fault" in res.stderr + res.stdout


def test_many_packed_workers_per_rank(tmp_path):
    """The ``large_scale`` shape in miniature: 24 workers as threads behind two ranks (12 per process), FedOBD with
    its two phases and quantized uploads.  Both replicas must finish and agree; the same run with 100 workers
    (50 threads per process) completes as well but is too slow for the suite."""
    conf = {**_TINY, "distributed_algorithm": "fed_obd", "worker_number": 24, "round": 2, "log_level": "WARNING",
            "dataset_kwargs": {"train_size": 24 * 64, "test_size": 160},
            "endpoint_kwargs": {"server": {"weight": 0.01}, "worker": {"weight": 0.01}},
            "algorithm_kwargs": {"dropout_rate": 0.3, "random_client_number": 24, "second_phase_epoch": 1}}
    script = tmp_path / "packed.py"
    script.write_text(MULTI_SCRIPT.format(root=ROOT, confs={"fed_obd": conf}, out=str(tmp_path)))
    env = dict(os.environ, DLS_FORCE_CPU="1", DLS_OUTPUT_ROOT=str(tmp_path), OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=420)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    r0, r1 = (json.load(open(tmp_path / f"multi_{r}.json")) for r in range(2))
    assert r0 == r1 and len(r0["fed_obd"]) >= 3      # two stage-one rounds and at least one stage-two record
