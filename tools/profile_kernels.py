#!/usr/bin/env python3
"""Launches every hand-written kernel a few times on realistic sizes (ResNet-18 sized flat
arena, ResNet/DenseNet activation shapes, LeNet-5 coalition GEMMs) and prints CUDA-event
timings with achieved bandwidth / FLOP rates.  Also the workload for `ncu --set full`."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_learning_simulator_b200.engine.model_util import ParameterLayout  # noqa: E402
from distributed_learning_simulator_b200.models.vision import ResNet18  # noqa: E402
from distributed_learning_simulator_b200.ops import bn as B  # noqa: E402
from distributed_learning_simulator_b200.ops import gemm as G  # noqa: E402
from distributed_learning_simulator_b200.ops import misc as M  # noqa: E402
from distributed_learning_simulator_b200.ops import quant as Q  # noqa: E402
from distributed_learning_simulator_b200.ops import reduce as R  # noqa: E402
from distributed_learning_simulator_b200.ops import sgd as S  # noqa: E402

from distributed_learning_simulator_b200.utils.clocks import ClockSampler  # noqa: E402

dev = torch.device("cuda", 0)
_clock_sampler = ClockSampler(gpu_index=0)
_clock_sampler.start()
REPS = int(os.environ.get("REPS", "20"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > L2 (126 MB)
results = {}


GRAPH_REPS = 10


def _replay_us(graph):
    graph.replay()
    torch.cuda.synchronize()
    samples = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        samples.append(e0.elapsed_time(e1) * 1e3)
    return sorted(samples)[len(samples) // 2]


_flush_graph_us = None


def _device_time(fn, cold):
    """Device time per call in us, free of the Python wrapper's host work: GRAPH_REPS calls captured in one CUDA
    graph (each preceded by an L2-evicting fill when ``cold``), minus the same graph with the fills alone."""
    global _flush_graph_us
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    with torch.cuda.graph(graph):
        for _ in range(GRAPH_REPS):
            if cold:
                flush.zero_()
            fn()
    total = _replay_us(graph)
    if cold:
        if _flush_graph_us is None:
            only = torch.cuda.CUDAGraph()
            with torch.cuda.graph(only):
                for _ in range(GRAPH_REPS):
                    flush.zero_()
            _flush_graph_us = _replay_us(only)
        total -= _flush_graph_us
    return total / GRAPH_REPS


def timed(name, fn, bytes_moved=None, flops=None, cold=True, capturable=True):
    """``us``: device time per call (CUDA-graph replay, no host work inside the measurement); ``us_eager``: the call
    as the framework's eager paths see it, Python wrapper included (events around one call after a synchronize)."""
    for _ in range(3):
        fn()
    times = []
    for _ in range(REPS):
        if cold:
            flush.zero_()  # evict L2 between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3)
    times.sort()
    eager = times[len(times) // 2]
    try:
        if not capturable:
            raise NotImplementedError("host work inside the call")
        us, how = _device_time(fn, cold), "graph"
    except Exception as e:  # noqa: BLE001 -- an op that cannot be captured keeps its eager number
        torch.cuda.synchronize()
        us, how = eager, f"eager ({type(e).__name__})"
    entry = {"us": round(us, 2), "us_eager": round(eager, 2), "timing": how}
    if bytes_moved:
        entry["GB/s"] = round(bytes_moved / us / 1e3, 1)
    if flops:
        entry["TFLOP/s"] = round(flops / us / 1e6, 2)
    results[name] = entry
    print(name, entry, flush=True)


model = ResNet18(3, 10)
layout = ParameterLayout.from_dict(dict(model.named_parameters()))
P = layout.total
rows = [torch.randn(P, device=dev) for _ in range(8)]
old = torch.randn(P, device=dev)
out = torch.empty(P, device=dev)
timed("fedavg_reduce_f32 K=8 P=11.2M", lambda: R.reduce_rows_f32(rows, [1.0] * 8, old, 1.0, 0.125, out=out, check_nan=False),
      bytes_moved=(8 + 2) * P * 4)
timed("fedavg_reduce_f32 K=2 P=11.2M", lambda: R.reduce_rows_f32(rows[:2], [1.0] * 2, old, 1.0, 0.5, out=out, check_nan=False),
      bytes_moved=(2 + 2) * P * 4)

flat, mom = torch.randn(P, device=dev), torch.zeros(P, device=dev)
grads = [torch.randn(n, device=dev) for n in layout.numels]
table = torch.tensor([g.data_ptr() for g in grads], dtype=torch.int64, device=dev)
offsets = layout.segment_offsets(dev)
lr = torch.full((8,), 0.01, device=dev)
state = torch.tensor([0, 1, 0, 0], dtype=torch.int32, device=dev)
timed("sgd_step momentum P=11.2M", lambda: S.sgd_step(flat, mom, table, offsets, len(layout.names), lr, state, 1e-4, 0.9),
      bytes_moved=5 * P * 4)

def sgd10():
    for _ in range(10):
        S.sgd_step(flat, mom, table, offsets, len(layout.names), lr, state, 1e-4, 0.9)


timed("sgd_step x10 back to back (working set 225 MB > L2)", sgd10, bytes_moved=10 * 5 * P * 4, cold=False)
rows8 = rows


def red10():
    for _ in range(10):
        R.reduce_rows_f32(rows8, [1.0] * 8, old, 1.0, 0.125, out=out, check_nan=False)


timed("fedavg_reduce_f32 K=8 x10 back to back", red10, bytes_moved=10 * (8 + 2) * P * 4, cold=False)
timed("sq_quant P=11.2M", lambda: Q.sq_quant(flat, layout, 255, seed=1), bytes_moved=P * (4 + 4 + 1 + 0.125))
q = Q.sq_quant(flat, layout, 255, seed=1)
deq = torch.empty(P, device=dev)
timed("sq_dequant P=11.2M", lambda: Q.sq_dequant(q, deq), bytes_moved=P * (1 + 0.125 + 4))
timed("nnadq_quant P=11.2M", lambda: Q.nnadq_quant(flat, layout, 0.001), bytes_moved=P * (4 + 4 + 1.2))
timed("sign_pack P=11.2M", lambda: M.sign_pack(flat), bytes_moved=P * (4 + 0.125))
packed = [M.sign_pack(r) for r in rows]
timed("sign_majority K=8", lambda: M.sign_majority(packed, P), bytes_moved=P * (8 * 0.125 + 4))
timed("dropout_mask P=11.2M", lambda: M.dropout_mask(flat, 0.7, seed=1), bytes_moved=P * 8)
block_of = torch.arange(len(layout.names), dtype=torch.int32, device=dev) // 3
n_blocks = int(block_of.max()) + 1
timed("block_sqdiff P=11.2M", lambda: M.block_sqdiff(flat, old, offsets, block_of, n_blocks),
      bytes_moved=P * 8)

for shape in [(64, 64, 32, 32), (64, 256, 8, 8), (64, 132, 32, 32)]:
    n, c, h, w = shape
    bn = torch.nn.BatchNorm2d(c, track_running_stats=False).to(dev)
    x = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    dy = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    elems = x.numel()
    with torch.no_grad():
        timed(f"bn_fwd+relu {shape}", lambda: B.bn_act(bn, x, relu=True), bytes_moved=elems * 2 * 2, cold=False)
    y = B.bn_act(bn, x, relu=True)
    timed(f"bn_bwd+relu {shape}", lambda: torch.autograd.grad(y, x, dy, retain_graph=True), bytes_moved=elems * 2 * 4,
          cold=False, capturable=False)

for (bsz, m, n, k) in [(1, 5000, 120, 400), (32, 5000, 120, 400), (32, 5000, 84, 120), (1, 8192, 4096, 4096)]:
    a = torch.randn(bsz, m, k, device=dev).bfloat16()
    w_ = torch.randn(bsz, n, k, device=dev).bfloat16()
    bias = torch.randn(bsz, n, device=dev)
    timed(f"gemm_tcgen05 b={bsz} {m}x{n}x{k}", lambda: G.linear_bf16(a, w_, bias, relu=True, out_dtype=torch.bfloat16),
          bytes_moved=bsz * (m * k + n * k + m * n) * 2, flops=2.0 * bsz * m * n * k, cold=False)

# coalition GEMM: W_s built inside the kernel (cluster pair sharing the tile / one CTA each) against
# the older two-step path (reduce kernel writes S models, cast, batched GEMM); LeNet-5 fc1 and conv2 shapes
from distributed_learning_simulator_b200.ops import runtime as RT  # noqa: E402

NW, S_, Pl = 10, 32, 61706
arenas = [torch.randn(Pl + 6, device=dev)[:Pl] for _ in range(NW)]
ptable = RT.pointer_table(arenas, dev)
member = torch.rand(S_, NW) < 0.5
member[:, 0] |= ~member.any(dim=1)
coefs = (member.float() / member.float().sum(1, keepdim=True)).to(dev).contiguous()
for (name, m, n, k, off) in [("fc1", 5000, 120, 400, 2572), ("conv2", 5000 * 100 // 10, 16, 150, 156)]:
    xa = torch.randn(S_, m, k + (-k) % 8, device=dev).bfloat16()
    for cl in (2, 1):
        os.environ["DLS_SUBSET_CLUSTER"] = str(cl)
        timed(f"subset_gemm {name} S={S_} NW={NW} {m}x{n}x{k} cluster={cl}",
              lambda: G.subset_linear(xa, arenas, ptable, coefs, off, off + n * k, n, k, relu=True,
                                      out_dtype=torch.bfloat16),
              flops=2.0 * S_ * m * n * k, cold=False, capturable=False)
    os.environ.pop("DLS_SUBSET_CLUSTER")
    stride = (Pl + 63) // 64 * 64
    models = torch.empty(S_, stride, device=dev)[:, :Pl]

    def two_step():
        for s_i in range(S_):
            sel = [i for i in range(NW) if member[s_i, i]]
            R.reduce_rows_f32([arenas[i] for i in sel], [1.0 / len(sel)] * len(sel), None, 0.0, 1.0, out=models[s_i],
                              check_nan=False)
        w_b = models[:, off: off + n * k].reshape(S_, n, k).to(torch.bfloat16)
        if k % 8:
            w_b = torch.nn.functional.pad(w_b, (0, 8 - k % 8))
        return G.linear_bf16(xa, w_b.contiguous(), models[:, off + n * k: off + n * k + n].contiguous(), relu=True,
                             out_dtype=torch.bfloat16)

    timed(f"materialise + gemm {name} S={S_} NW={NW} {m}x{n}x{k}", two_step, flops=2.0 * S_ * m * n * k, cold=False, capturable=False)

results["clocks"] = _clock_sampler.stop()  # SM clock / throttle record of this run
print("clocks", results["clocks"], flush=True)
out_path = os.environ.get("OUT", "gpurun_out/kernel_times.json")
os.makedirs(os.path.dirname(out_path), exist_ok=True)
with open(out_path, "w") as f:
    json.dump(results, f, indent=1)
