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
