"""csrc/graph.cu against plain PyTorch oracles (SURVEY K13 / K14 + the GCN aggregation)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture()
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from distributed_learning_simulator_b200.ops import runtime

    runtime.require()
    return torch.device("cuda", 0)


@pytest.mark.parametrize("n", [1, 31, 2047, 2048, 2049, 70001])
@pytest.mark.parametrize("through_n_id", [False, True])
def test_edge_compaction_is_stable_and_matches_boolean_indexing(dev, n, through_n_id):
    from distributed_learning_simulator_b200.ops import graph as G
    from distributed_learning_simulator_b200.ops import runtime

    g = torch.Generator().manual_seed(n)
    num_nodes, batch_nodes = 5000, 900
    mask = (torch.rand(num_nodes, generator=g) < 0.6).to(dev)
    if through_n_id:
        n_id = torch.randperm(num_nodes, generator=g)[:batch_nodes].to(dev)
        ei = torch.randint(0, batch_nodes, (2, n), generator=g).to(dev)
    else:
        n_id = None
        ei = torch.randint(0, num_nodes, (2, n), generator=g).to(dev)
    before = runtime.launch_count()
    got = G.filter_edges(ei, mask, n_id)
    assert runtime.launch_count() - before == 2
    keep = G.oracle_keep(ei[0], ei[1], mask, n_id)
    assert torch.equal(got, ei[:, keep])
    a, b, pos = G.compact_pairs(ei[0], ei[1], mask, n_id, want_pos=True)
    assert torch.equal(pos, keep.nonzero().view(-1)) and torch.equal(a, ei[0][keep]) and torch.equal(b, ei[1][keep])


def test_edge_drop_is_a_bernoulli_thinning_of_the_masked_edges(dev):
    from distributed_learning_simulator_b200.ops import graph as G

    g = torch.Generator().manual_seed(5)
    mask = (torch.rand(3000, generator=g) < 0.8).to(dev)
    ei = torch.randint(0, 3000, (2, 200000), generator=g).to(dev)
    full = G.oracle_keep(ei[0], ei[1], mask, None)
    a, b, pos = G.compact_pairs(ei[0], ei[1], mask, None, keep_prob=0.7, seed=99, want_pos=True)
    assert bool(full[pos].all()) and bool((pos[1:] > pos[:-1]).all())
    assert torch.equal(a, ei[0][pos]) and torch.equal(b, ei[1][pos])
    frac = pos.numel() / int(full.sum())
    assert abs(frac - 0.7) < 0.01, frac
    again = G.compact_pairs(ei[0], ei[1], mask, None, keep_prob=0.7, seed=99, want_pos=True)[2]
    other = G.compact_pairs(ei[0], ei[1], mask, None, keep_prob=0.7, seed=100, want_pos=True)[2]
    assert torch.equal(again, pos) and not (other.numel() == pos.numel() and torch.equal(other, pos))


@pytest.mark.parametrize("dtype,channels", [(torch.float32, 64), (torch.float32, 7), (torch.bfloat16, 6),
                                            (torch.bfloat16, 256)])
def test_boundary_select_and_row_gather(dev, dtype, channels):
    from distributed_learning_simulator_b200.ops import graph as G

    g = torch.Generator().manual_seed(channels)
    boundary = (torch.rand(4000, generator=g) < 0.3).to(dev)
    n_id = torch.randperm(4000, generator=g)[:1500].to(dev)
    x = torch.randn(1500, channels, generator=g).to(dev, dtype)
    rows, ids = G.select_nodes(n_id, boundary)
    want_rows = boundary[n_id].nonzero().view(-1)
    assert torch.equal(rows, want_rows) and torch.equal(ids, n_id[want_rows])
    assert torch.equal(G.gather_rows(x, rows), x.index_select(0, want_rows))


@pytest.mark.parametrize("channels", [64, 5])
@pytest.mark.parametrize("with_embedding", [True, False])
def test_halo_merge_forward_and_backward(dev, channels, with_embedding):
    from distributed_learning_simulator_b200.ops import graph as G

    g = torch.Generator().manual_seed(channels + int(with_embedding))
    train = (torch.rand(3000, generator=g) < 0.5).to(dev)
    n_id = torch.randperm(3000, generator=g)[:1200].to(dev)
    x = torch.randn(1200, channels, generator=g).to(dev).requires_grad_(True)
    x_ref = x.detach().clone().requires_grad_(True)
    ids = emb = None
    if with_embedding:
        # the server sends embeddings for some foreign nodes of the batch, some nodes outside it,
        # and (harmlessly) some local ones, in no particular order
        ids = torch.randperm(3000, generator=g)[:1700].to(dev)
        emb = torch.randn(1700, channels, generator=g).to(dev)
    out = G.halo_merge(x, n_id, train, ids, emb)
    ref = G.oracle_halo_merge(x_ref, n_id, train, ids, emb)
    assert torch.equal(out, ref)
    w = torch.randn(1200, channels, generator=g).to(dev)
    (out * w).sum().backward()
    (ref * w).sum().backward()
    assert torch.equal(x.grad, x_ref.grad)
    foreign = ~train[n_id]
    assert bool((x.grad[foreign] == 0).all())


@pytest.mark.parametrize("channels", [64, 7, 16])
def test_gcn_aggregate_forward_and_backward(dev, channels):
    from distributed_learning_simulator_b200.ops import graph as G

    g = torch.Generator().manual_seed(channels)
    n, e = 2000, 15000
    ei = torch.randint(0, n, (2, e), generator=g).to(dev)
    h = torch.randn(n, channels, generator=g).to(dev).requires_grad_(True)
    bias = torch.randn(channels, generator=g).to(dev).requires_grad_(True)
    h_ref, b_ref = h.detach().clone().requires_grad_(True), bias.detach().clone().requires_grad_(True)
    out = G.gcn_aggregate(h, ei, bias)
    ref = G.oracle_gcn_aggregate(h_ref.double(), ei, b_ref.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-5)
    w = torch.randn(n, channels, generator=g).to(dev)
    (out * w).sum().backward()
    (ref * w.double()).sum().backward()
    torch.testing.assert_close(h.grad, h_ref.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(bias.grad, b_ref.grad, rtol=1e-4, atol=1e-4)
    # no edges: only the self loop remains
    empty = torch.zeros(2, 0, dtype=torch.int64, device=dev)
    torch.testing.assert_close(G.gcn_aggregate(h.detach(), empty, None), h.detach())


def test_fed_gnn_trains_through_the_graph_kernels(dev):
    """Embedding-sharing fed_gnn on the GPU: per-batch edge filtering, boundary gather, halo merge and
    the GCN aggregation all run as this library's kernels (launch counter), and the run reaches the
    accuracy of the same configuration forced onto the PyTorch oracles."""
    from distributed_learning_simulator_b200 import load_config_from_file, train
    from distributed_learning_simulator_b200.ops import runtime
    from distributed_learning_simulator_b200.training import get_last_training_result

    overrides = {"round": 5, "epoch": 3, "worker_number": 3, "learning_rate": 0.05,
                 "dataset_kwargs": {"num_nodes": 600, "features": 32, "classes": 4},
                 "algorithm_kwargs": {"share_feature": True, "batch_number": 2, "num_neighbor": 5}}
    count0 = runtime.launch_count()
    config = load_config_from_file("fed_gnn/cs.yaml", overrides=overrides)
    config.round_timeout = 300
    train(config)
    res = get_last_training_result()
    assert sorted(res["performance"]) == [1, 2, 3, 4, 5]
    launched = runtime.launch_count() - count0
    assert launched > 200, launched
    acc = res["performance"][5]["test_accuracy"]
    assert acc > 0.3, acc
