#!/usr/bin/env python3
"""bf16 GEMM throughput: the hand-written tcgen05 kernels (single CTA per tile, CTA pair with
cta_group::2) and cuBLAS (torch.matmul) on the same shapes.  CUDA events, warm, median of 20."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200.ops import gemm as G  # noqa: E402

dev = torch.device("cuda", 0)


def timed(fn, reps=20):
    for _ in range(5):
        fn()
    times = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3)
    return sorted(times)[len(times) // 2]


rows = []
for (m, n, k) in [(8192, 4096, 4096), (8192, 8192, 8192), (4096, 4096, 1024), (16384, 1024, 1024), (2048, 2048, 2048)]:
    a = torch.randn(m, k, device=dev).bfloat16()
    w = torch.randn(n, k, device=dev).bfloat16()
    out = {"m": m, "n": n, "k": k}
    flops = 2.0 * m * n * k
    for name, env in (("pair", "1"), ("single", "0")):
        os.environ["DLS_GEMM_2CTA"] = env
        us = timed(lambda: G.linear_bf16(a, w, None, out_dtype=torch.bfloat16))
        out[f"{name}_us"] = round(us, 1)
        out[f"{name}_tflops"] = round(flops / us / 1e6, 1)
    us = timed(lambda: torch.matmul(a, w.t()))
    out["cublas_us"] = round(us, 1)
    out["cublas_tflops"] = round(flops / us / 1e6, 1)
    rows.append(out)
    print(json.dumps(out), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_bench.json", "w"), indent=1)
