#!/usr/bin/env python3
"""Aggregation bandwidth sweep (BASELINE.json config 5, second half): weighted FedAvg of one fp32
vector per rank, 1 MB - 1 GB, fused P2P reduce+broadcast kernel vs an NCCL all-reduce of
pre-weighted vectors (the baseline variant).  Launch with torch.distributed.run, one rank per GPU.

Timing: CUDA events on the launching stream, 3 warm-up + 10 timed iterations, max over ranks.
algbw = bytes of one vector / time; busbw = algbw * 2 (G-1)/G, the usual all-reduce convention.
Roofline of the fused kernel, per GPU and per NVLink direction: INGRESS = the peers' rows of my slice
arriving for my loads ((G-1)/G of a vector) + the other ranks' reduced slices arriving in my
global-model buffer ((G-1)/G of a vector) = 2 (G-1)/G * bytes; EGRESS is the mirror image with peer
stores (2 (G-1)/G * bytes) and drops to bytes with the NVLS multicast store (my slice leaves once, the
switch replicates it), so the bound is the ingress: time >= 2 (G-1)/G * bytes / (link bandwidth per
direction).  (An earlier version of this tool counted only one of the two ingress streams.)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from distributed_learning_simulator_b200.engine.model_util import ParameterLayout  # noqa: E402
from distributed_learning_simulator_b200.parallel.context import get_context  # noqa: E402

LINK_GBS = 770.0  # measured peer copy bandwidth per direction (B200_PROFILING.md); 900 nominal


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    times = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(times)[len(times) // 2]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ctx = get_context()
    G = ctx.world
    sizes_mb = [1, 4, 16, 64, 256, 1024]
    rows = []
    for mb in sizes_mb:
        n = mb * (1 << 20) // 4
        layout = ParameterLayout(["w"], [torch.Size([n])])
        ctx.setup_arena(layout, worker_number=G)
        flat = torch.randn(n, device=ctx.device)
        old = torch.zeros(n, device=ctx.device)
        ctx.store_upload(ctx.rank, flat)
        torch.cuda.synchronize()
        dist.barrier()
        payloads = [ctx.remote_payload(r, 0) for r in range(G)]
        weights = [1.0 + r for r in range(G)]
        inv = 1.0 / sum(weights)

        def fused():
            ctx.p2p_reduce(payloads, weights, old, 1.0, inv)

        pre = flat * (weights[ctx.rank] * inv)

        def nccl():
            buf = pre.clone()  # pre-weighting + copy is part of the NCCL variant's round
            dist.all_reduce(buf)
            buf.add_(old)

        # the broadcast half through NVSwitch multicast (one multimem.st per vector) when the platform has
        # it, and through one peer store per destination; the multicast-bound buffers serve both
        has_mc = ctx.multicast
        ctx.multicast_reduce = False
        ms_f = timed(fused)          # default: one peer store per destination
        ms_mc = None
        if has_mc:
            ctx.multicast_reduce = True
            ms_mc = timed(fused)     # broadcast half as multimem.st
            ctx.multicast_reduce = False
        ms_n = timed(nccl)
        out = ctx.p2p_reduce(payloads, weights, old, 1.0, inv).flat
        ref = pre.clone()
        dist.all_reduce(ref)
        err = float((out - ref).abs().max())
        bytes_ = n * 4
        roof_ms = 2 * (G - 1) / G * bytes_ / (LINK_GBS * 1e9) * 1e3
        rows.append({"MB": mb, "fused_ms": ms_f, "fused_multimem_store_ms": ms_mc, "nccl_ms": ms_n, "fused_algbw_GBs": bytes_ / ms_f / 1e6,
                     "nccl_algbw_GBs": bytes_ / ms_n / 1e6, "fused_busbw_GBs": bytes_ / ms_f / 1e6 * 2 * (G - 1) / G,
                     "roofline_ms": roof_ms, "fused_fraction_of_roofline": roof_ms / ms_f, "max_abs_err": err})
        if ctx.rank == 0:
            print(json.dumps(rows[-1]), flush=True)
    if ctx.rank == 0:
        out_path = os.environ.get("OUT", f"gpurun_out/agg_sweep_{G}gpu.json")
        os.makedirs(os.path.dirname(out_path), exist_ok=True)
        with open(out_path, "w") as f:
            json.dump({"gpus": G, "link_GBs": LINK_GBS, "rows": rows}, f, indent=1)
    dist.barrier()


if __name__ == "__main__":
    main()
