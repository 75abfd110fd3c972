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


def test_graph_ops_cpu_paths_agree_with_dense_formulas():
    """ops/graph.py on CPU (the oracles the CUDA kernels are tested against): edge filtering through
    n_id, boundary selection + row gather, halo merge with gradient masking, and the normalised GCN
    aggregation against the dense D^-1/2 (A + I) D^-1/2 h."""
    from distributed_learning_simulator_b200.ops import graph as G

    g = torch.Generator().manual_seed(0)
    num_nodes, batch_nodes, e = 50, 20, 200
    mask = torch.rand(num_nodes, generator=g) < 0.6
    n_id = torch.randperm(num_nodes, generator=g)[:batch_nodes]
    ei = torch.randint(0, batch_nodes, (2, e), generator=g)
    kept = G.filter_edges(ei, mask, n_id)
    want = mask[n_id[ei[0]]] & mask[n_id[ei[1]]]
    assert torch.equal(kept, ei[:, want])
    rows, ids = G.select_nodes(n_id, mask)
    assert torch.equal(rows, mask[n_id].nonzero().view(-1)) and torch.equal(ids, n_id[rows])
    x = torch.randn(batch_nodes, 8, generator=g, requires_grad=True)
    assert torch.equal(G.gather_rows(x, rows), x.detach()[rows])
    # halo merge: local rows pass through with gradient, foreign rows take the sent embedding or zero
    sent_ids = torch.cat([n_id[~mask[n_id]][:3], torch.tensor([num_nodes - 1])])
    emb = torch.randn(sent_ids.numel(), 8, generator=g)
    merged = G.halo_merge(x, n_id, mask, sent_ids, emb)
    for r in range(batch_nodes):
        node = int(n_id[r])
        if mask[node]:
            assert torch.equal(merged[r], x[r])
        elif node in sent_ids.tolist():
            assert torch.equal(merged[r], emb[sent_ids.tolist().index(node)])
        else:
            assert torch.equal(merged[r], torch.zeros(8))
    merged.sum().backward()
    assert torch.equal(x.grad, mask[n_id].float().unsqueeze(1).expand(-1, 8))
    # GCN aggregation against the dense operator
    n = 12
    edges = torch.randint(0, n, (2, 30), generator=g)
    h = torch.randn(n, 5, generator=g, dtype=torch.float64)
    bias = torch.randn(5, generator=g, dtype=torch.float64)
    adj = torch.zeros(n, n, dtype=torch.float64)
    for s, d in edges.t().tolist():
        adj[d, s] += 1.0
    adj += torch.eye(n, dtype=torch.float64)
    dinv = adj.sum(dim=1).rsqrt()
    dense = (dinv.unsqueeze(1) * adj * dinv.unsqueeze(0)) @ h + bias
    torch.testing.assert_close(G.gcn_aggregate(h, edges, bias), dense)
