#!/usr/bin/env python3
"""Micro-benchmark of one local training step (forward+backward, no optimizer) for layout /
precision variants, as CUDA-graph replays.  Guides the design of engine/graph_step.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200.models.vision import DenseNet40, ResNet18  # noqa: E402


def bench(name, model, x, y, autocast, steps=50):
    loss_fn = torch.nn.CrossEntropyLoss()
    params = list(model.parameters())

    def step():
        for p in params:
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            out = model(x)
            loss = loss_fn(out.float(), y)
        loss.backward()

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    eager = e0.elapsed_time(e1) / steps
    g = torch.cuda.CUDAGraph()
    for p in params:
        p.grad = None
    with torch.cuda.graph(g):
        step()
    torch.cuda.synchronize()
    for _ in range(5):
        g.replay()
    e0.record()
    for _ in range(steps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:40s} eager {eager:7.3f} ms   graph {e0.elapsed_time(e1) / steps:7.3f} ms", flush=True)


def main():
    torch.backends.cudnn.benchmark = os.environ.get("CUDNN_BENCH", "0") == "1"
    dev = torch.device("cuda")
    B = int(os.environ.get("BATCH", "64"))
    which = os.environ.get("MODEL", "resnet18")
    mk = (lambda: ResNet18(3, 10)) if which == "resnet18" else (lambda: DenseNet40(3, 10))
    x = torch.randn(B, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (B,), device=dev)
    only = os.environ.get("ONLY")
    variants = {
        "fp32 nchw": lambda: (mk().to(dev), x, False),
        "autocast bf16 nchw": lambda: (mk().to(dev), x, True),
        "autocast bf16 channels_last": lambda: (mk().to(dev).to(memory_format=torch.channels_last),
                                                x.contiguous(memory_format=torch.channels_last), True),
        "pure bf16 nchw": lambda: (mk().to(dev).bfloat16(), x.bfloat16(), False),
        "pure bf16 channels_last": lambda: (mk().to(dev).bfloat16().to(memory_format=torch.channels_last),
                                            x.bfloat16().contiguous(memory_format=torch.channels_last), False),
    }
    for name, make in variants.items():
        if only and only != name:
            continue
        model, xx, ac = make()
        for m in model.modules():
            if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
                m.track_running_stats = False
                m.register_buffer("running_mean", None)
                m.register_buffer("running_var", None)
                m.register_buffer("num_batches_tracked", None)
        model.train()
        bench(name, model, xx, y, ac)


if __name__ == "__main__":
    main()
