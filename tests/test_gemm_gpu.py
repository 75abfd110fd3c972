"""tcgen05/TMEM/TMA GEMM against a plain PyTorch fp32 matmul of the same bf16 operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (200, 10, 88), (5000, 120, 400), (37, 130, 1024), (1, 84, 120),
                                   (8192, 700, 256), (300, 300, 264)])
@pytest.mark.parametrize("bias,relu", [(False, False), (True, True)])
def test_linear_bf16(m, n, k, bias, relu):
    from distributed_learning_simulator_b200.ops import gemm as G
    from distributed_learning_simulator_b200.ops import runtime

    runtime.require()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn(m, k, device=dev, generator=g).bfloat16()
    w = (torch.randn(n, k, device=dev, generator=g) * 0.1).bfloat16()
    b = torch.randn(n, device=dev, generator=g) if bias else None
    out = G.linear_bf16(x, w, b, relu=relu)
    torch.cuda.synchronize()
    ref = G.oracle_linear(x, w, b, relu=relu)
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3 * (k ** 0.5))


@pytest.mark.parametrize("m,n,k,batch", [(1000, 300, 512, 1), (256, 128, 64, 1), (777, 1000, 264, 1), (513, 200, 136, 3),
                                         (4096, 2048, 1024, 1)])
@pytest.mark.parametrize("pair", ["1", "0"])
def test_linear_bf16_cta_pair(m, n, k, batch, pair, monkeypatch):
    """cta_group::2 variant (two CTAs on one 256 x N tile, leader-issued M = 256 MMAs) against the same
    oracle, including ragged M / N tails and batches; pair=0 pins the single-CTA kernel on the same shapes."""
    from distributed_learning_simulator_b200.ops import gemm as G

    monkeypatch.setenv("DLS_GEMM_2CTA", pair)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    x = torch.randn(batch, m, k, device=dev, generator=g).bfloat16()
    w = (torch.randn(batch, n, k, device=dev, generator=g) * 0.1).bfloat16()
    b = torch.randn(batch, n, device=dev, generator=g)
    out = G.linear_bf16(x, w, b, relu=True, out_dtype=torch.float32)
    torch.cuda.synchronize()
    ref = G.oracle_linear(x, w, b, relu=True)
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3 * (k ** 0.5))


def test_linear_bf16_batched_and_bf16_out():
    from distributed_learning_simulator_b200.ops import gemm as G

    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(5, 300, 152, device=dev, generator=g).bfloat16()
    w = (torch.randn(5, 70, 152, device=dev, generator=g) * 0.1).bfloat16()
    b = torch.randn(5, 70, device=dev, generator=g)
    out = G.linear_bf16(x, w, b, relu=True, out_dtype=torch.bfloat16)
    ref = G.oracle_linear(x, w, b, relu=True)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
    shared = G.linear_bf16(x, w[0], b[0])
    torch.testing.assert_close(shared, G.oracle_linear(x, w[0], b[0]), rtol=1e-3, atol=2e-2)
    # back-to-back launches reuse TMEM columns and barriers correctly
    again = G.linear_bf16(x, w[0], b[0])
    assert torch.equal(shared, again)


def test_inference_linear_patch_matches_torch():
    from distributed_learning_simulator_b200.ops import gemm as G

    dev = torch.device("cuda", 0)
    model = torch.nn.Sequential(torch.nn.Linear(512, 100), torch.nn.ReLU(), torch.nn.Linear(100, 10)).to(dev)
    x = torch.randn(1250, 512, device=dev)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        ref = model(x).float()
    assert G.install_inference_linear(model) == 2
    before = G.STATS["linear_calls"]
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        out = model(x).float()
    assert G.STATS["linear_calls"] - before == 2
    torch.testing.assert_close(out, ref, rtol=2e-2, atol=2e-2)
    assert (out.argmax(1) == ref.argmax(1)).float().mean() > 0.99
    y = model(x)  # autograd path is untouched
    y.sum().backward()
    assert model[0].weight.grad is not None and G.STATS["linear_calls"] - before == 2


@pytest.mark.parametrize("case", [
    # (coalitions, workers, M, n_layer, k_w, weight offset, stacked, bias, relu, out dtype, arena shift)
    dict(S=5, NW=10, M=300, n=120, k=400, off=2572, stacked=False, bias=True, relu=True, out=torch.bfloat16),
    dict(S=7, NW=10, M=1000, n=6, k=25, off=0, stacked=True, bias=True, relu=True, out=torch.bfloat16),
    dict(S=3, NW=4, M=257, n=16, k=150, off=156, stacked=False, bias=True, relu=False, out=torch.float32),
    dict(S=33, NW=10, M=64, n=10, k=84, off=60856, stacked=False, bias=False, relu=False, out=torch.float32),
    dict(S=4, NW=3, M=129, n=200, k=136, off=8, stacked=False, bias=True, relu=True, out=torch.float32),
    dict(S=9, NW=64, M=500, n=40, k=72, off=4, stacked=True, bias=True, relu=False, out=torch.float32),
    dict(S=2, NW=5, M=130, n=70, k=64, off=3, stacked=False, bias=True, relu=False, out=torch.float32, shift=1),
    dict(S=3, NW=6, M=5000, n=120, k=400, off=12, stacked=False, bias=True, relu=True, out=torch.float32),
    dict(S=2, NW=4, M=700, n=48, k=600, off=16, stacked=False, bias=True, relu=False, out=torch.float32),  # K > 512
    dict(S=6, NW=8, M=40000, n=6, k=25, off=0, stacked=True, bias=True, relu=True, out=torch.bfloat16),
])
@pytest.mark.parametrize("stationary", [1, 0])
@pytest.mark.parametrize("cluster", [2, 1])
def test_subset_gemm_builds_the_coalition_weights_inside_the_gemm(case, cluster, stationary, monkeypatch):
    """csrc/subset_gemm.cu: the B operand of the tcgen05 GEMM is formed in shared memory as the
    coalition-weighted sum of the workers' fp32 layer slices (never materialised).  Oracle: fp32
    einsum of the same slices, rounded to bf16 like the kernel's operand, then a plain matmul."""
    from distributed_learning_simulator_b200.ops import gemm, runtime

    runtime.require()
    # 2: CTA pairs share each built tile through DSMEM (default); 1: every CTA builds its own
    monkeypatch.setenv("DLS_SUBSET_CLUSTER", str(cluster))
    # 1: the built tile stays resident over a run of M tiles (K <= 512); 0: rebuilt for every output tile
    monkeypatch.setenv("DLS_SUBSET_STATIONARY", str(stationary))
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(case["S"] * 131 + case["n"])
    S, NW, M, n, k = case["S"], case["NW"], case["M"], case["n"], case["k"]
    total = case["off"] + n * k + n + 13
    shift = case.get("shift", 0)  # arenas that are only 4-byte aligned take the scalar load path
    arenas = [torch.randn(total + 4, generator=g).to(dev)[shift: shift + total] for _ in range(NW)]
    bias_off = case["off"] + n * k if case["bias"] else None
    member = torch.rand(S, NW, generator=g) < 0.5
    member[:, 0] |= ~member.any(dim=1)
    coefs = torch.rand(S, NW, generator=g) * member
    coefs = (coefs / coefs.sum(dim=1, keepdim=True)).to(dev).contiguous()
    shared_a = case["stacked"] or S % 2 == 1
    x = (torch.randn((M, k) if shared_a else (S, M, k), generator=g) * 0.5).to(dev, torch.bfloat16)
    table = runtime.pointer_table(arenas, dev)
    before = runtime.launch_count()
    out = gemm.subset_linear(x, arenas, table, coefs, case["off"], bias_off, n, k, stacked=case["stacked"],
                             relu=case["relu"], out_dtype=case["out"])
    torch.cuda.synchronize()
    assert runtime.launch_count() - before == 1
    ref = gemm.oracle_subset_linear(x, arenas, coefs, case["off"], bias_off, n, k, stacked=case["stacked"],
                                    relu=case["relu"])
    assert out.shape == ref.shape and out.dtype == case["out"]
    tol = 2e-2 if case["out"] == torch.bfloat16 else 2e-3
    torch.testing.assert_close(out.float(), ref, rtol=tol, atol=tol)
