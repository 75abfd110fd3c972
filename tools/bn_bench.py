#!/usr/bin/env python3
"""Times the fused BN(+ReLU) forward/backward against ATen's batch_norm + relu on the ResNet-18
layer shapes (channels-last bf16), as CUDA-graph replays."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200.ops import bn as B  # noqa: E402


INNER = 20


def timed(fn, steps=50):
    for _ in range(3):
        fn()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(INNER):  # many layers per graph: amortise the graph launch itself
            fn()
    for _ in range(5):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1000 / INNER


dev = torch.device("cuda")
for shape in [(64, 64, 32, 32), (64, 128, 16, 16), (64, 256, 8, 8), (64, 512, 4, 4), (64, 132, 32, 32)]:
    n, c, h, w = shape
    bn = torch.nn.BatchNorm2d(c, track_running_stats=False).to(dev)
    x = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    dy = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)

    def fused():
        x.grad = None
        bn.zero_grad(set_to_none=True)
        B.bn_act(bn, x, relu=True).backward(dy)

    def aten():
        x.grad = None
        bn.zero_grad(set_to_none=True)
        torch.relu(bn(x)).backward(dy)

    os.environ["DLS_FUSED_BN"] = "1"
    tf = timed(fused)
    os.environ["DLS_FUSED_BN"] = "0"
    ta = timed(aten)
    mb = x.numel() * 2 / 1e6
    print(f"{shape}: fused fwd+bwd {tf:7.1f} us   aten {ta:7.1f} us   ({mb:.1f} MB activations)", flush=True)
