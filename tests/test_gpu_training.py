"""GPU integration: the CUDA-graph training path and the fused aggregation inside a real
federated run, checked against the eager/torch path."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _conf(**kw):
    base = {"dataset_name": "MNIST", "model_name": "LeNet5", "distributed_algorithm": "fed_avg",
            "optimizer_name": "SGD", "learning_rate_scheduler_name": "CosineAnnealingLR", "worker_number": 2,
            "batch_size": 32, "round": 2, "epoch": 2, "learning_rate": 0.05,
            "dataset_kwargs": {"train_size": 700, "test_size": 200}}
    base.update(kw)
    return base


def test_graph_step_matches_eager():
    """Same data order, same init: the captured step (fused multi-tensor SGD, device-side LR
    schedule, remainder batch) reproduces the eager torch.optim.SGD loop.  Runs in a fresh
    process (tools/check_graph_step.py) with TF32 off."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, os.path.join(root, "tools", "check_graph_step.py")], capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    out = json.loads(res.stdout.strip().splitlines()[-1])
    assert out["replays"]["g1"] == 2 * 22 and out["replays"]["e1"] == 0
    assert max(out["diffs"].values()) < 1e-5, out["diffs"]
    for e in range(2):
        assert abs(out["metrics"]["g1"][e]["loss"] - out["metrics"]["e1"][e]["loss"]) < 1e-5
        assert abs(out["metrics"]["g1"][e]["accuracy"] - out["metrics"]["e1"][e]["accuracy"]) < 1e-6


def test_amp_graph_step_with_bf16_shadow_matches_autocast_eager():
    """use_amp: the captured step trains matrices/filters as bf16 views of the shadow arena (no cast
    kernels, the fused SGD kernel consumes bf16 gradients and rewrites the shadow); the eager arm is
    stock autocast + torch.optim.SGD.  The arithmetic is the same - bf16(master) weights, bf16
    gradients widened exactly - so with the same data order and init the fp32 masters agree."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    arms = [({}, 2 * 22), ({"DATASET": "CIFAR10", "MODEL": "resnet18", "TRAIN": "300", "EPOCHS": "1"}, 10)]
    for extra, replays in arms:
        env = dict(os.environ, AMP="1", ORDER="g1,e1", **extra)
        res = subprocess.run([sys.executable, os.path.join(root, "tools", "check_graph_step.py")],
                             capture_output=True, text=True, timeout=600, env=env)
        assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
        out = json.loads(res.stdout.strip().splitlines()[-1])
        assert out["replays"]["g1"] == replays and out["shadowed"]["g1"] >= 2 and out["shadowed"]["e1"] == 0
        # Both arms run the convolutions on the own kernels, but the captured step keeps the weight gradients of
        # the split layers in fp32 sinks while eager autocast rounds them to bf16 before torch.optim.SGD: the
        # masters agree to bf16 rounding of the updates, not bit for bit
        tol = 5e-2
        assert out["diffs"]["g1-e1"] < tol, out["diffs"]
        for g, e in zip(out["metrics"]["g1"], out["metrics"]["e1"]):
            assert abs(g["loss"] - e["loss"]) < 10 * tol and abs(g["accuracy"] - e["accuracy"]) < 10 * tol


def test_fed_avg_on_gpu_uses_fused_paths():
    from distributed_learning_simulator_b200 import load_config_from_dict, train
    from distributed_learning_simulator_b200.ops import runtime
    from distributed_learning_simulator_b200.training import get_last_training_result

    from distributed_learning_simulator_b200.engine import graph_step
    from distributed_learning_simulator_b200.ops import reduce as R

    before = runtime.launch_count()
    fused_before, replay_before = R.STATS["f32"], graph_step.REPLAY_COUNTER[0]
    config = load_config_from_dict(_conf(worker_number=3, learning_rate=0.1))
    train(config)
    perf = get_last_training_result()["performance"]
    assert sorted(perf) == [1, 2]
    assert perf[2]["test_accuracy"] > 0.15
    assert runtime.launch_count() > before
    assert R.STATS["f32"] - fused_before == 2, "each round must be aggregated by the fused kernel"
    assert graph_step.REPLAY_COUNTER[0] > replay_before, "local training must run as CUDA-graph replays"


def test_smoke_entry():
    import __graft_entry__ as entry

    entry.smoke()


def test_batched_coalition_evaluator_matches_sequential():
    """GTG-Shapley on GPU: coalitions evaluated in batched launches - the coalition GEMM that builds
    W_s inside the tcgen05 kernel (default), or the older materialise-then-GEMM path - give the same
    Shapley values as one-by-one evaluation, within the bf16 tolerance."""
    from distributed_learning_simulator_b200 import load_config_from_dict, train
    from distributed_learning_simulator_b200.ops import subset_eval
    from distributed_learning_simulator_b200.training import get_last_training_result

    results = {}
    modes = {"fused": {"batched_subset_eval": True}, "materialised": {"batched_subset_eval": True,
                                                                       "subset_eval_fused": False},
             "sequential": {"batched_subset_eval": False}}
    for mode, kwargs in modes.items():
        before = dict(subset_eval.STATS)
        config = load_config_from_dict(_conf(distributed_algorithm="GTG_shapley_value", worker_number=4, round=2,
                                             epoch=1, learning_rate=0.1, algorithm_kwargs=kwargs,
                                             dataset_kwargs={"train_size": 1200, "test_size": 2000}))
        train(config)
        results[mode] = get_last_training_result()
        used = subset_eval.STATS["coalitions"] - before["coalitions"]
        fused_launches = subset_eval.STATS["fused_layer_launches"] - before["fused_layer_launches"]
        assert (used > 0) == (mode != "sequential")
        assert (fused_launches > 0) == (mode == "fused")
    for mode in ("fused", "materialised"):
        for rnd in (1, 2):
            a, b = results[mode]["sv"][rnd], results["sequential"]["sv"][rnd]
            assert set(a) == set(b) == {0, 1, 2, 3}
            for w in a:
                assert abs(a[w] - b[w]) < 0.02, (mode, rnd, a, b)
        for rnd in (0, 1, 2):
            assert abs(results[mode]["performance"][rnd]["test_accuracy"]
                       - results["sequential"]["performance"][rnd]["test_accuracy"]) < 0.02  # independent trainings


def test_split_backward_convolution_matches_autograd():
    """ops/conv.py: weight gradients computed on a side stream (joined by the caller) match stock autograd,
    with and without autocast, for strided / padded / bias-carrying channels-last convolutions."""
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    side = torch.cuda.Stream(device=dev)
    g = torch.Generator(device="cuda").manual_seed(0)
    for amp, (cin, cout, k, stride, pad, bias) in [(False, (3, 16, 3, 1, 1, False)), (True, (16, 32, 3, 2, 1, False)),
                                                   (True, (32, 64, 1, 2, 0, True)), (False, (8, 8, 5, 1, 0, True))]:
        ref = torch.nn.Conv2d(cin, cout, k, stride=stride, padding=pad, bias=bias).to(dev).to(memory_format=torch.channels_last)
        new = torch.nn.Conv2d(cin, cout, k, stride=stride, padding=pad, bias=bias).to(dev).to(memory_format=torch.channels_last)
        new.load_state_dict(ref.state_dict())
        assert C.install(new) == 1
        x = torch.randn(8, cin, 16, 16, device=dev, generator=g).contiguous(memory_format=torch.channels_last)
        xs = [x.clone().requires_grad_(True) for _ in range(2)]
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            y_ref = ref(xs[0])
            with C.split_backward(side):
                y_new = new(xs[1])
        # own kernels (tensor-core / stem / direct route) against ATen: same dtype, equal to rounding
        assert y_new.dtype == y_ref.dtype
        torch.testing.assert_close(y_new.float(), y_ref.float(), rtol=2e-2, atol=2e-2 if amp else 5e-3)
        w = torch.randn_like(y_ref)
        (y_ref * w).sum().backward()
        (y_new * w).sum().backward()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        # different kernels on the two sides: compare to rounding rather than bit for bit
        tol = dict(rtol=2e-2, atol=2e-1) if amp else dict(rtol=1e-2, atol=1e-1)  # fp32 convs may run as TF32
        torch.testing.assert_close(xs[1].grad, xs[0].grad, **tol)
        assert new.weight.grad.dtype == new.weight.dtype
        torch.testing.assert_close(new.weight.grad, ref.weight.grad, **tol)
        if bias:
            torch.testing.assert_close(new.bias.grad, ref.bias.grad, **tol)
    # outside a split_backward context the layer takes the same route, on one stream
    before = sum(C.STATS.values())
    y = new(xs[1].detach())
    assert sum(C.STATS.values()) == before + 1
    torch.testing.assert_close(y.float(), y_ref.float(), rtol=2e-2, atol=2e-2)
