"""Every hand-written kernel against a plain PyTorch fp32/fp64 statement of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from distributed_learning_simulator_b200.ops import runtime

    runtime.require()
    return torch.device("cuda", 0)


def _layout(numels):
    from distributed_learning_simulator_b200.engine.model_util import ParameterLayout

    return ParameterLayout([f"t{i}" for i in range(len(numels))], [torch.Size([n]) for n in numels])


NUMELS = [1728, 64, 64, 36864, 10, 7, 1, 5120, 33, 4099]


@pytest.mark.parametrize("k", [1, 3, 8, 17])
@pytest.mark.parametrize("delta", [True, False])
def test_fedavg_reduce_f32(dev, k, delta):
    from distributed_learning_simulator_b200.ops import reduce as R

    n = sum(NUMELS) + 3
    g = torch.Generator(device="cuda").manual_seed(k)
    rows = [torch.randn(n, device=dev, generator=g) for _ in range(k)]
    weights = [float(100 + 7 * i) for i in range(k)]
    old = torch.randn(n, device=dev, generator=g)
    out = R.reduce_rows_f32(rows, weights, old if delta else None, 1.0 if delta else 0.0, 1.0 / sum(weights))
    ref = R.oracle_reduce(rows, weights, old if delta else None, 1.0 if delta else 0.0, 1.0 / sum(weights))
    torch.testing.assert_close(out, ref, rtol=1e-6, atol=1e-6)


def test_fedavg_reduce_nan_flag(dev):
    from distributed_learning_simulator_b200.ops import reduce as R

    rows = [torch.randn(1000, device=dev) for _ in range(2)]
    rows[1][17] = float("nan")
    with pytest.raises(AssertionError):
        R.reduce_rows_f32(rows, [1.0, 1.0], None, 0.0, 0.5)


def test_sq_quant_roundtrip_and_unbiased(dev):
    from distributed_learning_simulator_b200.ops import quant as Q

    layout = _layout(NUMELS)
    flat = torch.randn(layout.total, device=dev)
    flat[layout.offsets[2]:layout.offsets[3]] *= 1e-3
    flat[5] = 0.0
    q = Q.sq_quant(flat, layout, 255, seed=1)
    deq = Q.sq_dequant(q)
    offs = layout.offsets
    for t in range(len(NUMELS)):
        x = flat[offs[t]:offs[t + 1]]
        s = x.abs().max()
        torch.testing.assert_close(q.scale[t] * 255, s, rtol=1e-6, atol=0)
        assert (deq[offs[t]:offs[t + 1]] - x).abs().max() <= s / 255 * 1.0001 + 1e-12
    assert deq[5] == 0
    assert torch.equal(torch.sign(deq)[deq != 0], torch.sign(flat)[deq != 0])
    # unbiased: average of many independent quantizations converges to x
    acc = torch.zeros_like(flat)
    reps = 64
    for r in range(reps):
        acc += Q.sq_dequant(Q.sq_quant(flat, layout, 255, seed=100 + r))
    err = (acc / reps - flat).abs()
    step = torch.cat([q.scale[t].expand(NUMELS[t]) for t in range(len(NUMELS))])
    assert (err <= step * 0.5).float().mean() > 0.99
    # torch oracle dequant understands the kernel's packing
    from distributed_learning_simulator_b200.algorithm.quantization.packed import unpack_bits

    neg = unpack_bits(q.aux.cpu(), 1, layout.total).bool()
    assert torch.equal(neg.to(dev)[flat != 0], (flat < 0)[flat != 0])


@pytest.mark.parametrize("weight", [0.01, 0.001, 0.0001])
def test_nnadq_matches_oracle(dev, weight, monkeypatch):
    from distributed_learning_simulator_b200.algorithm.quantization import deterministic as D
    from distributed_learning_simulator_b200.ops import quant as Q

    layout = _layout(NUMELS)
    flat = torch.randn(layout.total, device=dev) * 0.05
    flat[layout.offsets[5]:layout.offsets[6]] = 0.25  # constant tensor: zero range
    q = Q.nnadq_quant(flat, layout, weight)
    deq = Q.nnadq_dequant(q)
    monkeypatch.setattr(Q, "available", lambda: False)
    ref_q = D.nnadq_quantize_flat(flat.cpu(), layout, weight)
    ref = D.nnadq_dequantize_flat(ref_q)
    assert torch.equal(q.bits.cpu(), ref_q.bits)
    assert torch.equal(q.byte_offsets.cpu(), ref_q.byte_offsets)
    torch.testing.assert_close(q.scale.cpu(), ref_q.scale, rtol=1e-6, atol=0)
    nbytes = int(ref_q.byte_offsets[-1])
    same = (q.payload[:nbytes].cpu() == ref_q.payload).float().mean()
    assert same > 0.999  # rounding ties may differ in the last ulp of the statistics
    step = torch.cat([ref_q.scale[t].expand(NUMELS[t]) for t in range(len(NUMELS))])
    assert ((deq.cpu() - ref).abs() <= step * 1.001 + 1e-7).all()
    assert ((deq.cpu() - flat.cpu()).abs() <= step * 0.5001 + 1e-7).all()
    assert q.compression_ratio() == pytest.approx(ref_q.compression_ratio(), rel=1e-6)


@pytest.mark.parametrize("variant", ["fedavg", "dropout"])
def test_fedavg_reduce_packed_mixed(dev, variant):
    from distributed_learning_simulator_b200.message import DeltaParameterMessage, FlatParameter, ParameterMessage
    from distributed_learning_simulator_b200.ops import quant as Q
    from distributed_learning_simulator_b200.ops import reduce as R

    layout = _layout(NUMELS)
    old = FlatParameter(torch.randn(layout.total, device=dev), layout)
    rows, msgs = [], {}
    for k in range(5):
        x = torch.randn(layout.total, device=dev) * 0.1
        if variant == "dropout":
            x = x * (torch.rand_like(x) < 0.7)
            msgs[k] = ParameterMessage(dataset_size=10 + k, parameter=FlatParameter(x, layout))
            rows.append((x, False))
        elif k % 3 == 0:
            msgs[k] = DeltaParameterMessage(dataset_size=10 + k, delta_parameter=FlatParameter(x, layout))
            rows.append((x, True))
        elif k % 3 == 1:
            q = Q.sq_quant(x, layout, 255, seed=k)
            msgs[k] = DeltaParameterMessage(dataset_size=10 + k, delta_parameter={"__quantized__": q})
            rows.append((Q.sq_dequant(q), True))
        else:
            q = Q.nnadq_quant(x, layout, 0.001)
            msgs[k] = ParameterMessage(dataset_size=10 + k, parameter={"__quantized__": q})
            rows.append((Q.nnadq_dequant(q), False))
    for m in msgs.values():
        assert R.can_fuse(m, old)
    out = R.fused_aggregate(msgs, old, variant=variant).flat
    acc = torch.zeros(layout.total, dtype=torch.float64, device=dev)
    den = torch.zeros(layout.total, dtype=torch.float64, device=dev)
    for k, (x, is_delta) in enumerate(rows):
        v = (old.flat + x if is_delta else x).double()
        w = torch.full_like(den, float(10 + k))
        if variant == "dropout":
            w = w * (v != 0)
        acc += w * v
        den += w
    den[den == 0] = 1
    torch.testing.assert_close(out, (acc / den).float(), rtol=1e-5, atol=1e-6)


def test_fedavg_reduce_packed_partial_blocks(dev):
    """FedOBD stage 1: dropped blocks contribute the old value (ParameterMessage.complete)."""
    from distributed_learning_simulator_b200.message import FlatParameter, ParameterMessage
    from distributed_learning_simulator_b200.ops import reduce as R

    layout = _layout(NUMELS)
    old = FlatParameter(torch.randn(layout.total, device=dev), layout)
    msgs, dense = {}, []
    for k in range(3):
        x = torch.randn(layout.total, device=dev)
        names = [n for i, n in enumerate(layout.names) if (i + k) % 2 == 0]
        msgs[k] = ParameterMessage(dataset_size=5 + k, parameter=R.PartialFlat(x, layout, names))
        full = old.flat.clone()
        for n in names:
            full[layout.slice_of(n)] = x[layout.slice_of(n)]
        dense.append(full)
        assert R.can_fuse(msgs[k], old)
    out = R.fused_aggregate(msgs, old, variant="fedavg").flat
    ref = R.oracle_reduce(dense, [5.0, 6.0, 7.0], None, 0.0, 1.0 / 18.0)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("momentum,nesterov,wd", [(0.9, False, 0.0), (0.9, True, 1e-4), (0.0, False, 5e-4)])
def test_sgd_step_matches_torch(dev, momentum, nesterov, wd):
    from distributed_learning_simulator_b200.ops import sgd as S

    layout = _layout(NUMELS)
    params = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in NUMELS]
    ref_opt = torch.optim.SGD(params, lr=0.1, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    flat = torch.cat([p.detach() for p in params]).clone()
    mom = torch.zeros_like(flat)
    offsets = torch.tensor(layout.offsets, dtype=torch.int64, device=dev)
    lr_table = torch.tensor([0.1, 0.05, 0.025], device=dev)
    state = torch.zeros(4, dtype=torch.int32, device=dev)
    for step in range(3):
        grads = [torch.randn(n, device=dev) for n in NUMELS]
        if step == 1:
            grads[3] = None  # a parameter without gradient is left untouched
        for p, g in zip(params, grads):
            p.grad = g
        for group in ref_opt.param_groups:
            group["lr"] = float(lr_table[step])
        ref_opt.step()
        table = torch.tensor([0 if g is None else g.data_ptr() for g in grads], dtype=torch.int64, device=dev)
        S.sgd_step(flat, mom, table, offsets, len(NUMELS), lr_table, state, wd, momentum, 0.0, nesterov)
        S.sgd_tick(state)
        torch.cuda.synchronize()
        ref = torch.cat([p.detach() for p in params])
        torch.testing.assert_close(flat, ref, rtol=1e-6, atol=1e-6)
    assert state.tolist()[:3] == [3, 1, 3]


def test_sgd_step_bf16_gradients_and_shadow(dev):
    """Mixed precision: flagged tensors hand the kernel a bf16 gradient and get their updated value
    written to the bf16 shadow arena; the fp32 master follows torch.optim.SGD on the widened gradient."""
    from distributed_learning_simulator_b200.ops import sgd as S

    layout = _layout(NUMELS)
    flags_host = [int(n >= 1000 or n == 33) for n in NUMELS]  # includes an odd-sized, unaligned segment
    params = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in NUMELS]
    ref_opt = torch.optim.SGD(params, lr=0.1, momentum=0.9, weight_decay=1e-4)
    flat = torch.cat([p.detach() for p in params]).clone()
    shadow = flat.to(torch.bfloat16)
    untouched = shadow.clone()
    mom = torch.zeros_like(flat)
    offsets = torch.tensor(layout.offsets, dtype=torch.int64, device=dev)
    flags = torch.tensor(flags_host, dtype=torch.uint8, device=dev)
    lr_table = torch.tensor([0.1, 0.05], device=dev)
    state = torch.zeros(4, dtype=torch.int32, device=dev)
    for step in range(2):
        grads = [torch.randn(n, device=dev).to(torch.bfloat16 if f else torch.float32)
                 for n, f in zip(NUMELS, flags_host)]
        for p, g in zip(params, grads):
            p.grad = g.float()
        for group in ref_opt.param_groups:
            group["lr"] = float(lr_table[step])
        ref_opt.step()
        table = torch.tensor([g.data_ptr() for g in grads], dtype=torch.int64, device=dev)
        S.sgd_step(flat, mom, table, offsets, len(NUMELS), lr_table, state, 1e-4, 0.9, 0.0, False,
                   grad_is_bf16=flags, shadow=shadow)
        S.sgd_tick(state)
        torch.cuda.synchronize()
        ref = torch.cat([p.detach() for p in params])
        torch.testing.assert_close(flat, ref, rtol=1e-6, atol=1e-6)
        for i, f in enumerate(flags_host):
            seg = slice(layout.offsets[i], layout.offsets[i] + NUMELS[i])
            if f:
                assert torch.equal(shadow[seg], flat[seg].to(torch.bfloat16))
            else:
                assert torch.equal(shadow[seg], untouched[seg])  # fp32 tensors have no shadow copy to keep


def test_gather_and_metric(dev):
    from distributed_learning_simulator_b200.ops import sgd as S

    x_all = torch.randn(200, 3, 8, 8, device=dev)
    y_all = torch.randint(0, 10, (200,), device=dev)
    order = torch.randperm(150, device=dev)
    state = torch.zeros(4, dtype=torch.int32, device=dev)
    state[2] = 2
    x = torch.empty(32, 3, 8, 8, device=dev)
    y = torch.empty(32, dtype=torch.int64, device=dev)
    S.gather_batch(x_all, y_all, order, 150, state, 2, x, y)
    idx = order[64:96]
    assert torch.equal(x, x_all[idx]) and torch.equal(y, y_all[idx])
    # the last, smaller batch of an epoch starts at index * regular batch size
    state[2] = 4
    xs = torch.empty(22, 3, 8, 8, device=dev)
    ys = torch.empty(22, dtype=torch.int64, device=dev)
    S.gather_batch(x_all, y_all, order, 150, state, 2, xs, ys, batch_stride=32)
    assert torch.equal(xs, x_all[order[128:150]]) and torch.equal(ys, y_all[order[128:150]])
    # channels-last destination: rows come out H,W,C
    xc = torch.empty(32, 3, 8, 8, device=dev).contiguous(memory_format=torch.channels_last)
    state[2] = 2
    S.gather_batch(x_all, y_all, order, 150, state, 2, xc, y)
    assert torch.equal(xc, x_all[idx]) and xc.is_contiguous(memory_format=torch.channels_last)
    logits = torch.randn(32, 10, device=dev)
    loss = torch.nn.functional.cross_entropy(logits, y).reshape(1)
    acc = torch.zeros(3, dtype=torch.float64, device=dev)
    S.metric_accumulate(logits, y, loss, acc)
    S.metric_accumulate(logits, y, loss, acc)
    correct = (logits.argmax(1) == y).sum().item()
    torch.testing.assert_close(acc.cpu(), torch.tensor([2 * loss.item() * 32, 2.0 * correct, 64.0], dtype=torch.float64))


def test_sign_pack_and_majority(dev):
    from distributed_learning_simulator_b200.ops import misc as M

    n = 100003
    grads = [torch.randn(n, device=dev) for _ in range(7)]
    packed = [M.sign_pack(g) for g in grads]
    for g, p in zip(grads, packed):
        assert torch.equal(p.cpu(), M.oracle_sign_pack(g.cpu()))
    vote = M.sign_majority(packed, n)
    ref = torch.sign(sum(torch.where(g < 0, -1.0, 1.0) for g in grads))
    assert torch.equal(vote, ref)
    assert torch.equal(M.oracle_sign_majority([p.cpu() for p in packed], n), ref.cpu())


def test_dropout_mask(dev):
    from distributed_learning_simulator_b200.ops import misc as M

    x = torch.randn(1_000_003, device=dev)
    out, cnt = M.dropout_mask(x, 0.7, seed=3)
    kept = out != 0
    assert torch.equal(out[kept], x[kept])
    assert int(cnt.item()) == int(kept.sum().item())
    assert abs(kept.float().mean().item() - 0.7) < 0.005
    out2, _ = M.dropout_mask(x, 0.7, seed=3)
    assert torch.equal(out, out2)


def test_block_sqdiff(dev):
    from distributed_learning_simulator_b200.ops import misc as M

    layout = _layout(NUMELS)
    cur = torch.randn(layout.total, device=dev)
    prev = torch.randn(layout.total, device=dev)
    offsets = torch.tensor(layout.offsets, dtype=torch.int64, device=dev)
    block_of = torch.tensor([0, 0, 1, 1, 2, 2, 2, 3, 4, 4], dtype=torch.int32, device=dev)
    got = M.block_sqdiff(cur, prev, offsets, block_of, 5)
    ref = M.oracle_block_sqdiff(cur, prev, offsets, block_of, 5)
    torch.testing.assert_close(got, ref, rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(64, 64, 32, 32), (7, 24, 5, 3), (3, 512, 4, 4), (16, 36, 8, 8), (16, 36, 32, 32)])
@pytest.mark.parametrize("relu,with_res", [(True, False), (True, True), (False, False)])
@pytest.mark.parametrize("arena", [False, True])
def test_fused_bn_act_matches_pytorch(dev, dtype, shape, relu, with_res, arena):
    """``arena``: the forward runs inside a per-step statistics arena, so the row-major shapes take the backward
    that adds its partial sums into a zeroed arena slice (dls_bn_bwd_acc) instead of folding them over a cluster."""
    import contextlib

    from distributed_learning_simulator_b200.ops import bn as B
    from distributed_learning_simulator_b200.ops import conv as C

    n, c, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(c + h)
    bn = torch.nn.BatchNorm2d(c, track_running_stats=False).to(dev)
    with torch.no_grad():
        bn.weight.copy_(torch.rand(c, device=dev, generator=g) + 0.5)
        bn.bias.copy_(torch.randn(c, device=dev, generator=g) * 0.1)
    x = (torch.randn(shape, device=dev, generator=g) * 1.5 + 0.3).to(dtype).contiguous(memory_format=torch.channels_last)
    res = torch.randn(shape, device=dev, generator=g).to(dtype).contiguous(memory_format=torch.channels_last) \
        if with_res else None
    dy = torch.randn(shape, device=dev, generator=g).to(dtype).contiguous(memory_format=torch.channels_last)
    x1 = x.clone().requires_grad_(True)
    r1 = res.clone().requires_grad_(True) if with_res else None
    assert B.can_fuse(bn, x1, r1)
    acc_before = B.STATS["bwd_acc"]
    slab = torch.zeros(4 * c, dtype=torch.float32, device=dev)
    with (C.stats_arena(slab) if arena else contextlib.nullcontext()):
        y = B.bn_act(bn, x1, relu=relu, residual=r1)
    y.backward(dy)
    row_major = n * h * w >= 8192 and c // (8 if dtype == torch.bfloat16 and c % 8 == 0 else 4) <= 128
    assert (B.STATS["bwd_acc"] == acc_before + 1) == (arena and row_major)
    got = (y, x1.grad, bn.weight.grad.clone(), bn.bias.grad.clone(), r1.grad if with_res else None)
    bn.zero_grad()
    x2 = x.float().clone().requires_grad_(True)
    r2 = res.float().clone().requires_grad_(True) if with_res else None
    pre = B.oracle_bn_act(x2, bn.weight, bn.bias, bn.eps, False, r2)
    y_ref = torch.relu(pre) if relu else pre
    # backward through the kernel's own ReLU mask: outputs within rounding of zero may differ
    masked = pre * (y.detach().float() > 0).float() if relu else pre
    masked.backward(dy.float())
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(got[0].float(), y_ref.detach(), **tol)
    torch.testing.assert_close(got[1].float(), x2.grad, **(dict(rtol=5e-2, atol=5e-2) if dtype == torch.bfloat16 else tol))
    scale = float(n * h * w) ** 0.5
    gtol = dict(rtol=2e-2, atol=2e-2 * scale) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-3 * scale)
    torch.testing.assert_close(got[2], bn.weight.grad, **gtol)
    torch.testing.assert_close(got[3], bn.bias.grad, **gtol)
    if with_res:
        torch.testing.assert_close(got[4].float(), r2.grad, **tol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("groups,channels,k,channels_last_k", [(5, 6, 5, True), (3, 6, 5, False), (4, 3, 3, True),
                                                               (1, 1, 5, False)])
def test_im2col_grouped_matches_unfold(dev, dtype, groups, channels, k, channels_last_k):
    from distributed_learning_simulator_b200.ops import misc as M

    x = torch.randn(7, 14, 14, groups * channels, device=dev).to(dtype)
    for pad in (0, 2, 1):  # 0: the plain kernel; > 0: the zero-padding variant (LeNet-5's first layer uses 2)
        got = M.im2col_grouped(x, groups, k, channels_last_k, pad=pad)
        ref = M.oracle_im2col_grouped(x, groups, k, channels_last_k, got.shape[-1], pad)
        assert got.shape == ref.shape and got.shape[-1] % 8 == 0
        assert torch.equal(got, ref), pad
