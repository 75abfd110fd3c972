"""Two ranks, two GPUs: the peer-mapped arena + fused P2P reduce/broadcast kernel against a
plain fp64 PyTorch reduction, then a short federated run through the public API.
Skipped unless two CUDA devices are visible (run with ``gpurun --gpus 2``)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import json, os, sys
sys.path.insert(0, {root!r})
import torch
from distributed_learning_simulator_b200.parallel.context import get_context
from distributed_learning_simulator_b200.engine.model_util import ParameterLayout

ctx = get_context()
assert ctx.use_p2p
numels = [1728, 64, 36864, 10, 7, 5120, 33, 4099, 1000003]
layout = ParameterLayout([f"t{{i}}" for i in range(len(numels))], [torch.Size([n]) for n in numels])
ctx.setup_arena(layout, worker_number=2 * ctx.world + 1)
n = layout.total
errs = []
old = torch.randn(n, device=ctx.device, generator=torch.Generator(device="cuda").manual_seed(1))
for rnd in range(3):
    payloads, rows, weights = [], [], []
    n_workers = 2 * ctx.world + 1
    for w in range(n_workers):
        g = torch.Generator(device="cuda").manual_seed(100 * rnd + w)
        row = torch.randn(n, device=ctx.device, generator=g)   # every rank can recompute every row
        rows.append(row)
        weights.append(float(10 + w))
        if ctx.is_local(w):
            ctx.store_upload(w, row)
    torch.cuda.synchronize()
    for w in range(n_workers):
        payloads.append(ctx.remote_payload(ctx.owner_of(w), ctx.slot_of(w)))
    out = ctx.p2p_reduce(payloads, weights, old, 1.0, 1.0 / sum(weights)).flat
    ref = old.double()
    for r, wt in zip(rows, weights):
        ref = ref + r.double() * (wt / sum(weights))
    errs.append(float((out.double() - ref).abs().max()))
    old = out.clone()
bcast_err = None
if ctx.multicast:
    # K9 on its own: rank 0 stores through the multicast address, every rank finds the data in its buffer
    from distributed_learning_simulator_b200.ops import runtime
    ctx.fence()
    src = torch.arange(n, device=ctx.device, dtype=torch.float32) * 0.5 + 3.0
    if ctx.rank == 0:
        runtime.call("dls_multimem_bcast_f32", src.data_ptr(), ctx.globals.mc_ptr, n, runtime.stream_ptr(ctx.device))
    ctx.fence()
    bcast_err = float((ctx.globals.local_tensor(torch.float32, n) - src).abs().max())
    ctx.fence()
with open(os.path.join({out!r}, "p2p_%d.json" % ctx.rank), "w") as f:
    json.dump({{"errs": errs, "reductions": ctx.p2p_reductions, "multicast": ctx.multicast,
               "multicast_reductions": ctx.multicast_reductions, "bcast_err": bcast_err,
               "two_stage": ctx.two_stage_reductions}}, f)

from distributed_learning_simulator_b200 import load_config_from_dict, train
from distributed_learning_simulator_b200.training import get_last_training_result
conf = {{"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
        "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 2 * ctx.world,
        "batch_size": 32, "round": 3, "epoch": 1, "learning_rate": 0.1, "log_level": "WARNING",
        "dataset_kwargs": {{"train_size": 640 * ctx.world, "test_size": 400}}}}
before = ctx.p2p_reductions
train(load_config_from_dict(conf))
res = get_last_training_result()
with open(os.path.join({out!r}, "train_%d.json" % ctx.rank), "w") as f:
    json.dump({{"performance": {{str(k): v for k, v in res["performance"].items()}},
               "p2p_reductions": ctx.p2p_reductions - before}}, f)
"""


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("nvls,two_stage,world", [("1", "1", 2), ("0", "1", 2), ("0", "0", 2), ("0", "1", 4), ("1", "1", 8)])
def test_p2p_reduce_and_distributed_training(tmp_path, nvls, two_stage, world):
    """nvls=1: the broadcast half of the fused kernel is one multimem.st per vector (NVSwitch multicast)
    when the platform offers it; nvls=0: one peer store per destination.  two_stage=1: each rank folds its own
    2-3 rows into one partial before the cross-GPU kernel; 0: every row crosses NVLink.  Same results either way."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    script = tmp_path / "run.py"
    script.write_text(SCRIPT.format(root=ROOT, out=str(tmp_path)))
    env = dict(os.environ, DLS_OUTPUT_ROOT=str(tmp_path), DLS_NVLS=nvls, DLS_NVLS_REDUCE=nvls, DLS_TWO_STAGE=two_stage)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    for r in range(world):
        p2p = json.load(open(tmp_path / f"p2p_{r}.json"))
        assert max(p2p["errs"]) < 1e-5, p2p
        assert p2p["reductions"] == 3 and p2p["two_stage"] == (3 if two_stage == "1" else 0)
        if nvls == "0":
            assert not p2p["multicast"] and p2p["multicast_reductions"] == 0
        elif p2p["multicast"]:
            assert p2p["multicast_reductions"] == 3 and p2p["bcast_err"] == 0.0
        print("multicast available:", p2p["multicast"])
    runs = [json.load(open(tmp_path / f"train_{r}.json")) for r in range(world)]
    assert all(t["performance"] == runs[0]["performance"] and t["p2p_reductions"] == 3 for t in runs)
    assert runs[0]["performance"]["3"]["test_accuracy"] > 0.2


STRESS = r"""
import json, os, random, sys, time
sys.path.insert(0, {root!r})
import torch
from distributed_learning_simulator_b200.parallel.context import get_context
from distributed_learning_simulator_b200.engine.model_util import ParameterLayout

ctx = get_context()
layout = ParameterLayout(["a", "b"], [torch.Size([65536 + 123]), torch.Size([4099])])
ctx.setup_arena(layout, worker_number=3 * ctx.world)
n = layout.total
rng = random.Random(1234 + ctx.rank)          # per-rank jitter, same data on every rank (seeded by round)
noise_stream = torch.cuda.Stream(device=ctx.device)
a = torch.randn(4096, 4096, device=ctx.device, dtype=torch.bfloat16)
worst = 0.0
old = torch.zeros(n, device=ctx.device)
for rnd in range(200):
    with torch.cuda.stream(noise_stream):      # keeps the SMs busy while the reduction's CTAs spin on peer flags
        for _ in range(rng.randint(0, 3)):
            a @ a
    workers = 3 * ctx.world if rnd % 2 else ctx.world   # two-stage and direct rounds alternate
    rows, weights, payloads = [], [], []
    for w in range(workers):
        g = torch.Generator(device="cuda").manual_seed(1000 * rnd + w)
        row = torch.randn(n, device=ctx.device, generator=g)
        rows.append(row); weights.append(float(1 + (w * 7 + rnd) % 5))
        if ctx.is_local(w):
            ctx.store_upload(w, row)
    if rng.random() < 0.3:
        time.sleep(rng.random() * 0.002)       # ranks arrive at the kernel's flag barrier at different times
    for w in range(workers):
        payloads.append(ctx.remote_payload(ctx.owner_of(w), ctx.slot_of(w)))
    out = ctx.p2p_reduce(payloads, weights, old, 0.5, 1.0 / sum(weights)).flat
    ref = 0.5 * old.double()
    for r, wt in zip(rows, weights):
        ref += r.double() * (wt / sum(weights))
    worst = max(worst, float((out.double() - ref).abs().max()))
    old = out.clone()
ctx.finish_reductions()
torch.cuda.synchronize()
with open(os.path.join({out!r}, "stress_%d.json" % ctx.rank), "w") as f:
    json.dump({{"worst": worst, "reductions": ctx.p2p_reductions, "two_stage": ctx.two_stage_reductions}}, f)
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_p2p_reduce_flag_protocol_stress(tmp_path):
    """200 rounds on every visible GPU (up to 8): random arrival skew at the device-side flag barrier, SM-saturating
    kernels on a second stream of every rank while the reduction's CTAs wait on peer flags, alternating one and
    three rows per rank (direct and two-stage paths); every result is compared with an fp64 oracle."""
    world = min(torch.cuda.device_count(), 8)
    script = tmp_path / "stress.py"
    script.write_text(STRESS.format(root=ROOT, out=str(tmp_path)))
    env = dict(os.environ, DLS_OUTPUT_ROOT=str(tmp_path), DLS_NVLS="0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    for r in range(world):
        s = json.load(open(tmp_path / f"stress_{r}.json"))
        assert s["worst"] < 1e-5 and s["reductions"] == 200 and s["two_stage"] == 100, s
