"""Fused training-mode BatchNorm (+ residual add) (+ ReLU) on channels-last activations
(csrc/bn.cu): one cluster launch forward and one backward (statistics exchanged through
distributed shared memory), instead of ATen's separate statistics / transform / ReLU / add kernels.  ``bn_act`` is what the vision models call; it falls back to
the plain modules whenever the fused kernel does not apply (CPU, running statistics enabled,
NCHW activations, odd channel counts)."""
import os

import torch
from torch import nn

from . import runtime

_DTYPE = {torch.float32: 0, torch.bfloat16: 1}
STATS = {"fused_fwd": 0, "fused_stats": 0, "fallback": 0, "bwd_acc": 0}


def _enabled() -> bool:
    return os.environ.get("DLS_FUSED_BN", "1") != "0"


def can_fuse(bn: nn.Module, x: torch.Tensor, residual: torch.Tensor | None) -> bool:
    if not (_enabled() and x.is_cuda and x.dim() == 4 and x.dtype in _DTYPE and runtime.available()):
        return False
    if bn.track_running_stats or bn.running_mean is not None or not bn.affine:
        return False  # the kernel implements the batch-statistics mode this framework trains in
    c = x.shape[1]
    vec = 8 if (x.dtype == torch.bfloat16 and c % 8 == 0) else 4
    if c % 4 or c < 4:
        return False
    if not x.is_contiguous(memory_format=torch.channels_last):
        return False
    if residual is not None and (residual.dtype != x.dtype or residual.shape != x.shape
                                 or not residual.is_contiguous(memory_format=torch.channels_last)):
        return False
    return True


_BWD_ACC = os.environ.get("DLS_BN_BWD_ACC", "1") != "0"


def _scratch(rows: int, channels: int, dtype: int, device: torch.device) -> torch.Tensor | None:
    """Per-call partial-sum buffer of the row-major path (a fresh tensor: safe under graph capture)."""
    need = runtime.require().dls_bn_scratch_floats(rows, channels, dtype)
    return torch.empty(need, dtype=torch.float32, device=device) if need > 0 else None


class _FusedBNAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual, eps, relu, bn, stats=None):
        n, c, h, w = x.shape
        rows = n * h * w
        dtype = _DTYPE[x.dtype]
        y = torch.empty_like(x)  # preserves channels-last
        mean = torch.empty(c, dtype=torch.float32, device=x.device)
        invstd = torch.empty(c, dtype=torch.float32, device=x.device)
        if stats is not None:
            # the convolution that produced x left sum / sum of squares per channel in its epilogue
            runtime.call("dls_bn_fwd_stats", x.data_ptr(), dtype, rows, c, weight.data_ptr(), bias.data_ptr(),
                         runtime.ptr(residual), float(eps), int(relu), y.data_ptr(), mean.data_ptr(),
                         invstd.data_ptr(), stats.data_ptr(), runtime.stream_ptr(x.device))
            STATS["fused_stats"] += 1
        else:
            scratch = _scratch(rows, c, dtype, x.device)
            runtime.call("dls_bn_fwd", x.data_ptr(), dtype, rows, c, weight.data_ptr(), bias.data_ptr(),
                         runtime.ptr(residual), float(eps), int(relu), y.data_ptr(), mean.data_ptr(),
                         invstd.data_ptr(), runtime.ptr(scratch), runtime.stream_ptr(x.device))
        STATS["fused_fwd"] += 1
        # row-major shapes: a zeroed slice of the step's arena lets the backward reduction add its partial sums with
        # atomics instead of folding them across a cluster (csrc/bn.cu, dls_bn_bwd_acc)
        ctx.bwd_partials = None
        if _BWD_ACC and any(ctx.needs_input_grad[:3]) and runtime.require().dls_bn_scratch_floats(rows, c, dtype) > 0:
            from . import conv as conv_ops

            ctx.bwd_partials = conv_ops.take_arena_slice(2 * c)
        ctx.save_for_backward(x, y if relu else x, weight, mean, invstd)
        ctx.relu, ctx.has_residual, ctx.bn = bool(relu), residual is not None, bn
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, weight, mean, invstd = ctx.saved_tensors
        n, c, h, w = x.shape
        rows = n * h * w
        dtype = _DTYPE[x.dtype]
        if dy.dtype != x.dtype or not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.to(x.dtype).contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if ctx.has_residual else None
        dgamma = torch.empty(c, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(c, dtype=torch.float32, device=x.device)
        if ctx.bwd_partials is not None:
            runtime.call("dls_bn_bwd_acc", dy.data_ptr(), y.data_ptr(), x.data_ptr(), dtype, rows, c, weight.data_ptr(),
                         mean.data_ptr(), invstd.data_ptr(), int(ctx.relu), dx.data_ptr(), runtime.ptr(dres),
                         dgamma.data_ptr(), dbeta.data_ptr(), ctx.bwd_partials.data_ptr(),
                         runtime.stream_ptr(x.device))
            STATS["bwd_acc"] += 1
            ctx.bwd_partials = None  # the slice is dirty now: a second backward over this node takes the scratch path
            return dx, dgamma, dbeta, dres, None, None, None, None
        scratch = _scratch(rows, c, dtype, x.device)
        runtime.call("dls_bn_bwd", dy.data_ptr(), y.data_ptr(), x.data_ptr(), dtype, rows, c, weight.data_ptr(),
                     mean.data_ptr(), invstd.data_ptr(), int(ctx.relu), dx.data_ptr(), runtime.ptr(dres),
                     dgamma.data_ptr(), dbeta.data_ptr(), runtime.ptr(scratch), runtime.stream_ptr(x.device))
        return dx, dgamma, dbeta, dres, None, None, None, None


def bn_act(bn: nn.Module, x: torch.Tensor, relu: bool = False, residual: torch.Tensor | None = None,
           stats: torch.Tensor | None = None) -> torch.Tensor:
    """``relu(bn(x) + residual)`` with batch statistics.  ``stats`` ([2, C] fp32: per-channel sum and sum of
    squares of ``x``, e.g. from the epilogue of the convolution that produced it) skips the reduction pass."""
    if can_fuse(bn, x, residual):
        if stats is not None and (x.shape[1] // (8 if x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0 else 4)) > 512:
            stats = None
        return _FusedBNAct.apply(x, bn.weight, bn.bias, residual, bn.eps, relu, bn, stats)
    STATS["fallback"] += 1
    out = bn(x)
    if residual is not None:
        out = out + residual
    return torch.relu(out) if relu else out


def oracle_bn_act(x, weight, bias, eps, relu, residual=None):
    """Plain PyTorch fp32 statement of the same op (test oracle)."""
    xf = x.float()
    mean = xf.mean(dim=(0, 2, 3), keepdim=True)
    var = xf.var(dim=(0, 2, 3), unbiased=False, keepdim=True)
    out = (xf - mean) * torch.rsqrt(var + eps) * weight.view(1, -1, 1, 1) + bias.view(1, -1, 1, 1)
    if residual is not None:
        out = out + residual.float()
    return torch.relu(out) if relu else out


def conv_bn_act(conv: nn.Module, bn: nn.Module, x: torch.Tensor, relu: bool = False,
                residual: torch.Tensor | None = None) -> torch.Tensor:
    """``relu(bn(conv(x)) + residual)``: on the tensor-core convolution path the BatchNorm statistics come out
    of the convolution's epilogue, so the normalisation is a single elementwise launch."""
    from . import conv as conv_ops

    if getattr(conv, "_dls_own_conv", False) and os.environ.get("DLS_CONV_BN_STATS", "1") != "0" \
            and not (bn.track_running_stats or bn.running_mean is not None):
        y, stats = conv_ops.forward_with_stats(conv, x, want_stats=True)
        return bn_act(bn, y, relu=relu, residual=residual, stats=stats)
    return bn_act(bn, conv(x), relu=relu, residual=residual)
