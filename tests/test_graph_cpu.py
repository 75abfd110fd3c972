"""Federated GNN family on a small synthetic graph (reference smoke test: test.sh line 6)."""
import json
import os

import pytest
import torch

from distributed_learning_simulator_b200 import load_config_from_file, train
from distributed_learning_simulator_b200.training import get_last_training_result


@pytest.mark.parametrize("conf,kwargs", [
    ("fed_gnn/cs.yaml", {"share_feature": True, "batch_number": 2, "edge_drop_rate": 0.5, "num_neighbor": 5}),
    ("fed_gnn/yelp.yaml", {"share_feature": False, "batch_number": 2, "edge_drop_rate": 0.9}),  # fed_gcn
    ("fed_aas/cora.yaml", None),
])
def test_graph_methods(monkeypatch, conf, kwargs):
    monkeypatch.setenv("DLS_FORCE_CPU", "1")
    overrides = {"round": 2, "epoch": 1, "worker_number": 3, "learning_rate": 0.01,
                 "dataset_kwargs": {"num_nodes": 300, "features": 16, "classes": 4}}
    if kwargs is not None:
        overrides["algorithm_kwargs"] = kwargs
    config = load_config_from_file(conf, overrides=overrides)
    config.round_timeout = 120
    train(config)
    res = get_last_training_result()
    assert sorted(res["performance"]) == [1, 2]
    for w in range(3):
        with open(os.path.join(res["save_dir"], f"worker_{w}", "graph_worker_stat.json"), encoding="utf8") as f:
            stat = json.load(f)
        assert stat["training_node_cnt"] > 0 and "communicated_bytes" in stat
        if conf == "fed_gnn/cs.yaml":
            assert stat["in_client_training_edge_cnt"] <= stat["original_in_client_training_edge_cnt"]


def test_embedding_exchange_algorithm():
    from distributed_learning_simulator_b200.algorithm.graph_algorithm import GraphNodeEmbeddingPassingAlgorithm
    from distributed_learning_simulator_b200.message import Message

    algo = GraphNodeEmbeddingPassingAlgorithm()
    e0 = torch.arange(6, dtype=torch.float32).view(3, 2)
    algo.process_worker_data(0, Message(in_round=True, other_data={
        "node_embedding": (e0, torch.tensor([10, 11, 12])), "boundary": torch.tensor([20, 21])}), None, "")
    algo.process_worker_data(1, Message(in_round=True, other_data={
        "node_embedding": (e0 + 100, [20, 22, 23]), "boundary": {11, 12, 99}}), None, "")
    res = algo.aggregate_worker_data().other_data["worker_result"]
    assert res[0]["node_indices"] == (20,) and torch.equal(res[0]["node_embedding"], (e0 + 100)[:1])
    assert res[1]["node_indices"] == (11, 12) and torch.equal(res[1]["node_embedding"], e0[1:])
