"""Race evidence for the mbarrier / cluster protocols of the tensor-core kernels without compute-sanitizer:
many launches with randomised shapes while other streams keep every SM busy, every output compared with a
torch oracle.  A lost arrival, a stale parity or a ring slot overwritten too early shows up as a wrong tile or
as the kernels' bounded-wait trap."""
import random

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _noise(dev, stop_after: int):
    """SM-saturating background work on two extra streams (large matmuls + elementwise)."""
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    b = torch.randn(8 << 20, device=dev)
    for i in range(stop_after):
        with torch.cuda.stream(streams[i % 2]):
            if i % 2:
                a @ a
            else:
                b.mul_(1.0001).add_(1e-6)
    return streams


def test_conv_kernels_under_concurrency():
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    rng = random.Random(7)
    g = torch.Generator(device="cuda").manual_seed(7)
    launches = 0
    for it in range(60):
        n = rng.choice([1, 3, 8, 16, 64])
        ci = rng.choice([8, 24, 64, 128, 256, 512])
        co = rng.choice([8, 40, 64, 128, 256, 512])
        hw = rng.choice([4, 7, 8, 16, 32]) if n * ci <= 64 * 128 else rng.choice([4, 8])
        k, st = rng.choice([(1, 1), (3, 1), (3, 2), (1, 2)])
        if st == 2 and (hw % 2 or ci % 64):
            st = 1
        pad = k // 2
        x = torch.randn(n, ci, hw, hw, device=dev, generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, k, k, device=dev, generator=g) / (ci * k * k) ** 0.5).bfloat16().contiguous(
            memory_format=torch.channels_last)
        shape = C.tc_shape(tuple(x.shape), tuple(w.shape), (st, st), (pad, pad))
        if shape is None:
            continue
        xf, wf = x.float().requires_grad_(True), w.float().requires_grad_(True)
        yf = F.conv2d(xf, wf, None, st, pad)
        dy = torch.randn(yf.shape, device=dev, generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
        yf.backward(dy.float())
        torch.cuda.synchronize()
        streams = _noise(dev, 6)
        outs = []
        for _ in range(3):  # back to back: ring / accumulator state must not leak from launch to launch
            stats = torch.zeros(2, co, device=dev)
            outs.append((C.tc_fprop(x, w, shape, None, stats), C.tc_dgrad(dy, w, shape),
                         C.tc_wgrad(dy, x, shape, accumulate=True), stats))
            launches += 3
        for s in streams:
            s.synchronize()
        torch.cuda.synchronize()
        for y, dx, dw, stats in outs:
            for got, ref in ((y, yf.detach()), (dx, xf.grad), (dw, wf.grad)):
                err = float((got.float() - ref).norm() / ref.norm().clamp_min(1e-6))
                assert err < 1e-2, (it, tuple(x.shape), tuple(w.shape), st, err)
            assert torch.isfinite(stats).all()
    assert launches >= 300


def test_gemm_kernels_under_concurrency():
    from distributed_learning_simulator_b200.ops import gemm as G

    dev = torch.device("cuda", 0)
    rng = random.Random(11)
    g = torch.Generator(device="cuda").manual_seed(11)
    for it in range(120):
        m, n, k = rng.choice([1, 37, 128, 300, 1000, 4096]), rng.choice([10, 64, 84, 200, 512]), 8 * rng.randint(1, 96)
        x = torch.randn(m, k, device=dev, generator=g).bfloat16()
        w = (torch.randn(n, k, device=dev, generator=g) * 0.1).bfloat16()
        streams = _noise(dev, 4)
        out = G.linear_bf16(x, w, None)
        for s in streams:
            s.synchronize()
        torch.cuda.synchronize()
        ref = G.oracle_linear(x, w, None)
        torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3 * (k ** 0.5))
