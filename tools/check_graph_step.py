#!/usr/bin/env python3
"""Trains the same model on the same data order with the CUDA-graph step and with the eager
torch.optim.SGD loop and prints the largest parameter difference (JSON).  Used by
tests/test_gpu_training.py in a fresh process so nothing else has touched cuDNN state."""
import itertools
import json
import os
import sys

# cuDNN autotuning picks algorithms by timing, so two runs may round differently: pin the heuristics
os.environ.setdefault("DLS_CUDNN_BENCHMARK", "0")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200 import load_config_from_dict  # noqa: E402
from distributed_learning_simulator_b200.engine.ml_type import MachineLearningPhase  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
EPOCHS = int(os.environ.get("EPOCHS", "2"))
conf = {"dataset_name": os.environ.get("DATASET", "MNIST"), "model_name": os.environ.get("MODEL", "LeNet5"),
        "distributed_algorithm": "fed_avg", "optimizer_name": "SGD",
        "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 2, "batch_size": 32, "round": 2,
        "epoch": EPOCHS, "learning_rate": 0.05, "log_level": "WARNING", "weight_decay": 1e-4,
        "dataset_kwargs": {"train_size": int(os.environ.get("TRAIN", "700")), "test_size": 200}}
if os.environ.get("AMP", "0") == "1":
    conf["use_amp"] = True  # graph arm trains on the bf16 shadow arena, eager arm under plain autocast
ORDER = os.environ.get("ORDER", "g1,e1,g2").split(",")
params, metrics, replays, shadowed = {}, {}, {}, {}
for tag in ORDER:
    config = load_config_from_dict(dict(conf, use_cuda_graph=tag.startswith("g")))
    trainer = config.create_trainer()
    trainer.dataset_collection.remove_dataset(MachineLearningPhase.Validation)
    trainer.set_seed(123)
    trainer.set_device(torch.device("cuda", 0))
    trainer.train()
    trainer.wait_stream()
    params[tag] = trainer.model_util.get_parameter_list().clone()
    metrics[tag] = [trainer.performance_metric.get_epoch_metrics(e) for e in range(1, EPOCHS + 1)]
    replays[tag] = trainer._graph_step.replays if trainer._graph_step is not None else 0
    shadowed[tag] = sum(p.dtype == torch.bfloat16 for p in trainer.model.parameters())
diffs = {f"{a}-{b}": float((params[a] - params[b]).abs().max()) for a, b in itertools.combinations(ORDER, 2)}
print(json.dumps({"diffs": diffs, "metrics": metrics, "replays": replays, "shadowed": shadowed}))
