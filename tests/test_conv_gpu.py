"""Hand-written convolutions (tcgen05 implicit GEMM fprop / dgrad / wgrad, direct small-channel kernels)
against plain PyTorch fp32 ``conv2d`` + autograd on the same bf16-rounded operands, for every layer shape
of ResNet-18 (CIFAR stem), LeNet-5, and shapes with partial boxes / channel tails."""
import pytest
import torch
import torch.nn.functional as F
from torch import nn

pytestmark = pytest.mark.gpu

# (n, c_in, h, w, c_out, k, stride, pad)
RESNET18 = [
    (64, 64, 32, 32, 64, 3, 1, 1), (64, 64, 32, 32, 128, 3, 2, 1), (64, 64, 32, 32, 128, 1, 2, 0),
    (64, 128, 16, 16, 128, 3, 1, 1), (64, 128, 16, 16, 256, 3, 2, 1), (64, 128, 16, 16, 256, 1, 2, 0),
    (64, 256, 8, 8, 256, 3, 1, 1), (64, 256, 8, 8, 512, 3, 2, 1), (64, 256, 8, 8, 512, 1, 2, 0),
    (64, 512, 4, 4, 512, 3, 1, 1),
]
ODD = [(10, 24, 12, 12, 40, 3, 1, 1), (3, 72, 7, 9, 16, 1, 1, 0), (5, 64, 10, 6, 72, 3, 2, 1), (37, 8, 28, 28, 8, 5, 1, 2),
       (1250, 64, 32, 32, 64, 3, 1, 1),
       # halo-patch kernel: partial tile in h, channel tails on both sides, several filter sets
       (3, 40, 24, 16, 72, 3, 1, 1), (2, 136, 16, 24, 64, 3, 1, 1), (5, 64, 40, 32, 64, 3, 1, 1)]


def _case(n, ci, h, w, co, k, st, pad, seed=0):
    from distributed_learning_simulator_b200.ops import conv as C
    from distributed_learning_simulator_b200.ops import runtime

    runtime.require()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(seed + n + ci + co)
    x = torch.randn(n, ci, h, w, device=dev, generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(co, ci, k, k, device=dev, generator=g) / (ci * k * k) ** 0.5).bfloat16().contiguous(
        memory_format=torch.channels_last)
    shape = C.tc_shape(tuple(x.shape), tuple(wt.shape), (st, st), (pad, pad))
    assert shape is not None
    xf, wf = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yf = F.conv2d(xf, wf, None, st, pad)
    dy = torch.randn(yf.shape, device=dev, generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    yf.backward(dy.float())
    return C, x, wt, dy, shape, yf.detach(), xf.grad, wf.grad


def _close(out, ref, k_red):
    # bf16 operands, fp32 accumulation; outputs rounded to bf16 where the kernel stores bf16
    scale = float(ref.abs().max())
    tol = scale * (8e-3 if out.dtype == torch.bfloat16 else 1e-4 + 3e-7 * k_red ** 0.5) + 1e-6 * k_red ** 0.5
    torch.testing.assert_close(out.float(), ref, rtol=0, atol=tol)


@pytest.mark.parametrize("case", RESNET18 + ODD)
def test_fprop_and_stats(case):
    C, x, wt, dy, shape, y_ref, _, _ = _case(*case)
    stats = torch.zeros(2, case[4], device=x.device)
    bias = torch.randn(case[4], device=x.device)
    y = C.tc_fprop(x, wt, shape, None, stats)
    _close(y, y_ref, case[1] * case[5] ** 2)
    yb = y.float()
    torch.testing.assert_close(stats[0], yb.sum((0, 2, 3)), rtol=1e-3, atol=1e-3 * float(yb.abs().sum((0, 2, 3)).max()))
    torch.testing.assert_close(stats[1], (yb * yb).sum((0, 2, 3)), rtol=1e-3, atol=1e-3)
    y2 = C.tc_fprop(x, wt, shape, bias)
    _close(y2, y_ref + bias.view(1, -1, 1, 1), case[1] * case[5] ** 2)


@pytest.mark.parametrize("case", RESNET18 + ODD)
def test_dgrad(case):
    C, x, wt, dy, shape, _, dx_ref, _ = _case(*case)
    dx = C.tc_dgrad(dy, wt, shape)
    _close(dx, dx_ref, case[4] * case[5] ** 2)


@pytest.mark.parametrize("case", RESNET18 + ODD)
def test_wgrad(case):
    C, x, wt, dy, shape, _, _, dw_ref = _case(*case)
    dw = C.tc_wgrad(dy, x, shape, accumulate=True)
    _close(dw, dw_ref, case[0] * case[2] * case[3])
    dw2 = torch.full_like(dw, 3.0)  # store mode overwrites whatever is there
    C.tc_wgrad(dy, x, shape, out=dw2, accumulate=False)
    # one CTA walks the whole pixel range in store mode: a single fp32 chain of up to 1.3 M products
    torch.testing.assert_close(dw2, dw_ref, rtol=0, atol=float(dw_ref.abs().max()) * (1e-4 + 2e-6 * (case[0] * case[2] * case[3]) ** 0.5))
    dwh = torch.full_like(dw, 3.0).bfloat16()  # the same, stored as bf16 (gradient sinks of non-split layers)
    C.tc_wgrad(dy, x, shape, out=dwh, accumulate=False)
    torch.testing.assert_close(dwh.float(), dw_ref, rtol=0, atol=float(dw_ref.abs().max()) * (8e-3 + 2e-6 * (case[0] * case[2] * case[3]) ** 0.5))
    C.tc_wgrad(dy, x, shape, out=dw, accumulate=True)  # a second pass adds
    _close(dw, 2 * dw_ref, case[0] * case[2] * case[3])


@pytest.mark.parametrize("n,ci,h,w,co,k,st,pad,bias,amp", [
    (64, 3, 32, 32, 64, 3, 1, 1, False, False),    # ResNet-18 / DenseNet stem in fp32
    (32, 1, 28, 28, 6, 5, 1, 2, True, False),      # LeNet-5 conv1
    (32, 6, 14, 14, 16, 5, 1, 0, True, False),     # LeNet-5 conv2
    (8, 3, 17, 13, 10, 3, 2, 1, True, False),
])
def test_direct_module(n, ci, h, w, co, k, st, pad, bias, amp):
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    torch.manual_seed(n + ci + co)
    conv = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev)
    ref = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev).double()  # fp64 oracle (cuDNN's fp32 path is TF32)
    ref.load_state_dict(conv.state_dict())
    C.install(conv)
    x = torch.randn(n, ci, h, w, device=dev)
    if ci > 1:
        x = x.contiguous(memory_format=torch.channels_last)
    x1, x2 = x.clone().requires_grad_(True), x.double().requires_grad_(True)
    before = C.STATS["direct"]
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
        y = conv(x1)
    assert C.STATS["direct"] == before + 1, "the layer did not take the direct kernels"
    yr = ref(x2)
    g = torch.randn_like(yr)
    y.backward(g.to(y.dtype))
    yr.backward(g)
    tol = 2e-2 if amp else 1e-4
    torch.testing.assert_close(y.double(), yr, rtol=tol, atol=tol)
    torch.testing.assert_close(x1.grad.double(), x2.grad, rtol=tol, atol=tol * float(x2.grad.abs().max()))
    torch.testing.assert_close(conv.weight.grad.double(), ref.weight.grad, rtol=tol,
                               atol=tol * float(ref.weight.grad.abs().max()))
    if bias:
        torch.testing.assert_close(conv.bias.grad.double(), ref.bias.grad, rtol=tol, atol=tol * float(ref.bias.grad.abs().max()))


def test_module_path_matches_autocast_reference():
    """A small residual stack through ``install``: forward, input gradient and weight gradients of the
    tensor-core route against eager bf16 autocast on the ATen path."""
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    torch.manual_seed(5)

    def make():
        return nn.Sequential(nn.Conv2d(16, 64, 3, 1, 1, bias=False), nn.ReLU(), nn.Conv2d(64, 128, 3, 2, 1, bias=True),
                             nn.ReLU(), nn.Conv2d(128, 128, 1, 1, 0, bias=False)).to(dev).to(memory_format=torch.channels_last)

    own, ref = make(), make()
    ref.load_state_dict(own.state_dict())
    C.install(own)
    x = torch.randn(16, 16, 20, 20, device=dev).contiguous(memory_format=torch.channels_last)
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    before = C.STATS["tc"]
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y1 = own(x1)
        y2 = ref(x2)
    assert C.STATS["tc"] == before + 3
    g = torch.randn_like(y2)
    y1.backward(g)
    y2.backward(g)
    torch.testing.assert_close(y1.float(), y2.float(), rtol=3e-2, atol=3e-2)
    # two bf16 pipelines that round intermediates at different places: a few elements differ by ~1 bf16 ulp of the
    # largest intermediate
    torch.testing.assert_close(x1.grad, x2.grad, rtol=5e-2, atol=1e-1 * float(x2.grad.abs().max()))
    for (name, p1), (_, p2) in zip(own.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(p1.grad, p2.grad, rtol=5e-2, atol=1e-1 * float(p2.grad.abs().max()),
                                   msg=lambda m, name=name: f"{name}: {m}")


@pytest.mark.parametrize("n,ci,h,w,co,k,st,pad,bias", [(64, 3, 32, 32, 64, 3, 1, 1, False), (16, 3, 32, 32, 24, 3, 1, 1, True),
                                                       (9, 1, 28, 28, 16, 5, 1, 2, True), (4, 3, 30, 22, 64, 7, 2, 3, False)])
@pytest.mark.parametrize("channels_last_weight", [True, False])
def test_stem_module(n, ci, h, w, co, k, st, pad, bias, channels_last_weight):
    """Few-input-channel layers under bf16 autocast: explicit patch matrix + the tensor-core 1x1 kernels."""
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    torch.manual_seed(n + ci + co)
    conv = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev)
    ref = nn.Conv2d(ci, co, k, st, pad, bias=bias).to(dev)
    if channels_last_weight:
        conv = conv.to(memory_format=torch.channels_last)
    ref.load_state_dict(conv.state_dict())
    C.install(conv)
    x = torch.randn(n, ci, h, w, device=dev).contiguous(memory_format=torch.channels_last)
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    before = C.STATS["stem"]
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = conv(x1)
    assert C.STATS["stem"] == before + 1, "the layer did not take the stem route"
    yr = ref(x2)
    g = torch.randn_like(yr)
    y.backward(g.to(y.dtype))
    yr.backward(g)
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(x1.grad, x2.grad, rtol=2e-2, atol=2e-2 * float(x2.grad.abs().max()))
    torch.testing.assert_close(conv.weight.grad, ref.weight.grad, rtol=2e-2, atol=2e-2 * float(ref.weight.grad.abs().max()))
    if bias:
        torch.testing.assert_close(conv.bias.grad, ref.bias.grad, rtol=2e-2, atol=2e-2 * float(ref.bias.grad.abs().max()))


@pytest.mark.parametrize("n,ci,hw,co,st,relu,res", [(64, 64, 32, 64, 1, True, True), (64, 128, 16, 256, 2, False, False),
                                                    (64, 512, 4, 512, 1, True, False), (64, 3, 32, 64, 1, True, False)])
def test_conv_bn_act_statistics_from_epilogue(n, ci, hw, co, st, relu, res):
    """BatchNorm fed by the statistics of the convolution epilogue == BatchNorm that reduces its input itself,
    forward and backward."""
    from distributed_learning_simulator_b200.ops import bn as B
    from distributed_learning_simulator_b200.ops import conv as C

    dev = torch.device("cuda", 0)
    torch.manual_seed(ci + co)
    conv = nn.Conv2d(ci, co, 3, st, 1, bias=False).to(dev).to(memory_format=torch.channels_last)
    bn = nn.BatchNorm2d(co, track_running_stats=False).to(dev)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    C.install(conv)
    x = torch.randn(n, ci, hw, hw, device=dev).contiguous(memory_format=torch.channels_last)
    residual = torch.randn(n, co, hw // st, hw // st, device=dev).bfloat16().contiguous(memory_format=torch.channels_last) \
        if res else None
    outs = []
    for fused in (True, False):
        xi = x.clone().requires_grad_(True)
        for p in list(conv.parameters()) + list(bn.parameters()):
            p.grad = None
        before = B.STATS["fused_stats"]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            if fused:
                y = B.conv_bn_act(conv, bn, xi, relu=relu, residual=residual)
            else:
                y = B.bn_act(bn, conv(xi), relu=relu, residual=residual)
        assert (B.STATS["fused_stats"] == before + 1) == fused
        y.float().square().sum().backward()  # noqa
        outs.append((y.float(), xi.grad.clone(), conv.weight.grad.clone(), bn.weight.grad.clone(), bn.bias.grad.clone()))
    for a, b in zip(*outs):  # statistics summed in a different order: bf16-level differences that backward amplifies
        rel = float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))
        assert rel < 3e-2, rel
    # and against the plain fp32 statement of the op
    with torch.no_grad():
        ref = B.oracle_bn_act(torch.nn.functional.conv2d(x.bfloat16().float(), conv.weight.bfloat16().float(), None, st, 1)
                              .bfloat16(), bn.weight, bn.bias, bn.eps, relu, residual)
    torch.testing.assert_close(outs[0][0], ref, rtol=3e-2, atol=3e-2)
