#!/usr/bin/env python3
"""Numerics + timing table of the convolution kernels against a plain PyTorch fp32 oracle and the cuDNN
bf16 kernels, for every layer shape of ResNet-18 (CIFAR stem) at batch 64 plus a few odd shapes.

    python tools/conv_check.py [fprop|dgrad|wgrad|direct|all] [--time]

Each kernel kind runs in its own process (a faulting kernel must not take the others' results along).
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (n, c_in, h, w, c_out, k, stride, pad)
SHAPES = [
    (64, 64, 32, 32, 64, 3, 1, 1),
    (64, 64, 32, 32, 128, 3, 2, 1),
    (64, 64, 32, 32, 128, 1, 2, 0),
    (64, 128, 16, 16, 128, 3, 1, 1),
    (64, 128, 16, 16, 256, 3, 2, 1),
    (64, 128, 16, 16, 256, 1, 2, 0),
    (64, 256, 8, 8, 256, 3, 1, 1),
    (64, 256, 8, 8, 512, 3, 2, 1),
    (64, 256, 8, 8, 512, 1, 2, 0),
    (64, 512, 4, 4, 512, 3, 1, 1),
    (10, 24, 12, 12, 40, 3, 1, 1),     # odd sizes: partial boxes, channel tails
    (3, 72, 7, 9, 16, 1, 1, 0),
    (5, 64, 10, 6, 72, 3, 2, 1),
    (3, 40, 24, 16, 72, 3, 1, 1),      # halo-patch kernel with tails
    (2, 136, 16, 24, 64, 3, 1, 1),
]


def rel_err(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))


def timeit(fn, iters=20):
    """Device time per call in us: ``iters`` calls captured back to back in one CUDA graph (no host launch cost
    inside the measurement), median of 5 replays."""
    import torch

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(iters):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    times = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1) / iters * 1e3)
    return sorted(times)[2]


def timeit_eager(fn, iters=20):
    import torch

    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def run(kind: str, timing: bool) -> None:
    import torch
    import torch.nn.functional as F

    from distributed_learning_simulator_b200.ops import conv as C

    torch.manual_seed(0)
    dev = torch.device("cuda")
    rows = []
    for (n, ci, h, w, co, k, st, pad) in SHAPES:
        x = torch.randn(n, ci, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        wt = (torch.randn(co, ci, k, k, device=dev) / (ci * k * k) ** 0.5).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        shape = C.tc_shape(tuple(x.shape), tuple(wt.shape), (st, st), (pad, pad))
        tag = f"n{n} c{ci} {h}x{w} -> k{co} {k}x{k} s{st}"
        if shape is None:
            rows.append({"shape": tag, "kind": kind, "skipped": "not eligible"})
            continue
        xf, wf = x.float().requires_grad_(True), wt.float().requires_grad_(True)
        yf = F.conv2d(xf, wf, None, st, pad)
        dy = torch.randn_like(yf).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        yf.backward(dy.float())
        row = {"shape": tag, "kind": kind}
        if kind == "fprop":
            stats = torch.zeros(2, co, device=dev)
            y = C.tc_fprop(x, wt, shape, None, stats)
            row["err"] = rel_err(y, yf)
            yb = y.float()
            row["stats_err"] = max(rel_err(stats[0], yb.sum((0, 2, 3))), rel_err(stats[1], (yb * yb).sum((0, 2, 3))))
            if timing:
                row["us"] = timeit(lambda: C.tc_fprop(x, wt, shape))
                row["us_stats"] = timeit(lambda: C.tc_fprop(x, wt, shape, None, stats))
                row["cudnn_us"] = timeit(lambda: F.conv2d(x, wt, None, st, pad))
        elif kind == "dgrad":
            dx = C.tc_dgrad(dy, wt, shape)
            row["err"] = rel_err(dx, xf.grad)
            if timing:
                row["us"] = timeit(lambda: C.tc_dgrad(dy, wt, shape))
                row["cudnn_us"] = timeit(lambda: torch.ops.aten.convolution_backward(
                    dy, x, wt, None, [st, st], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False]))
        elif kind == "wgrad":
            dw = C.tc_wgrad(dy, x, shape, accumulate=True)
            row["err"] = rel_err(dw, wf.grad)
            dw2 = torch.full_like(dw, 7.0)
            C.tc_wgrad(dy, x, shape, out=dw2, accumulate=False)
            row["err_store"] = rel_err(dw2, wf.grad)
            if timing:
                row["us"] = timeit(lambda: C.tc_wgrad(dy, x, shape, out=dw, accumulate=True))
                row["us_store"] = timeit(lambda: C.tc_wgrad(dy, x, shape, out=dw2, accumulate=False))
                row["cudnn_us"] = timeit(lambda: torch.ops.aten.convolution_backward(
                    dy, x, wt, None, [st, st], [pad, pad], [1, 1], False, [0, 0], 1, [False, True, False]))
        torch.cuda.synchronize()
        rows.append(row)
        print(json.dumps(row), flush=True)


def run_direct(timing: bool) -> None:
    import torch
    import torch.nn.functional as F
    from torch import nn

    from distributed_learning_simulator_b200.ops import conv as C

    torch.manual_seed(0)
    dev = torch.device("cuda")
    cases = [(64, 3, 32, 32, 64, 3, 1, 1, False, True), (32, 1, 28, 28, 6, 5, 1, 2, True, False),
             (32, 6, 14, 14, 16, 5, 1, 0, True, False), (8, 3, 17, 13, 10, 3, 2, 1, True, False)]
    for (n, ci, h, w, co, k, st, pad, bias, amp) in cases:
        conv = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev)
        ref = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev)
        ref.load_state_dict(conv.state_dict())
        C.install(conv)
        x = torch.randn(n, ci, h, w, device=dev)
        if ci > 1:
            x = x.contiguous(memory_format=torch.channels_last)
        x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
        before = dict(C.STATS)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            y = conv(x1)
        yr = ref(x2)
        g = torch.randn_like(yr)
        y.backward(g.to(y.dtype))
        yr.backward(g)
        row = {"shape": f"direct n{n} c{ci} {h}x{w} -> k{co} {k}x{k} s{st} amp={amp}", "route": {
            key: C.STATS[key] - before[key] for key in C.STATS}, "err": rel_err(y, yr), "dx_err": rel_err(x1.grad, x2.grad),
            "dw_err": rel_err(conv.weight.grad, ref.weight.grad)}
        if bias:
            row["db_err"] = rel_err(conv.bias.grad, ref.bias.grad)
        if timing:
            def step(m, xx):
                xx.grad = None
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                    out = m(xx)
                out.backward(g.to(out.dtype))
            row["us_fwd_bwd_eager"] = timeit_eager(lambda: step(conv, x1))
            row["aten_us_fwd_bwd_eager"] = timeit_eager(lambda: step(ref, x2))
        print(json.dumps(row), flush=True)


def main() -> None:
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    timing = "--time" in sys.argv
    kind = args[0] if args else "all"
    if kind == "all":
        from distributed_learning_simulator_b200.utils.clocks import ClockSampler

        sampler = ClockSampler(gpu_index=0)
        sampler.start()
        for k in ("fprop", "dgrad", "wgrad", "direct"):
            cmd = [sys.executable, os.path.abspath(__file__), k] + (["--time"] if timing else [])
            res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
            print(f"==== {k}: rc {res.returncode}")
            print(res.stdout[-6000:])
            if res.returncode != 0:
                print(res.stderr[-3000:])
        print(json.dumps({"clocks": sampler.stop()}))
        return
    if kind == "direct":
        run_direct(timing)
    else:
        run(kind, timing)


if __name__ == "__main__":
    main()
