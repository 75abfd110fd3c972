"""``nn.Conv2d`` on this framework's own kernels (csrc/conv_tc.cu, csrc/conv_direct.cu).

The reference has no convolution code of its own: every convolution of a local step runs in cuDNN
inside the external eager ``Trainer.train()`` (/root/reference/simulation_lib/worker/worker.py:86-88)
and of a server evaluation inside ``tester.inference()`` (simulation_lib/server/server.py:40-55).

``install(model)`` re-routes the ``nn.Conv2d`` layers of a model through ``_Conv2d``:

* tensor-core path -- bf16 compute (autocast or bf16 inputs), channels-last activations, channel
  counts that are multiples of 8, stride 1 or 2: implicit-GEMM fprop / dgrad / wgrad on tcgen05,
  filters read in place from the channels-last arena segment (bf16 shadow), no im2col buffer, no
  transposed filter copy;
* direct path -- small filters (stems with 1 or 3 input channels, LeNet-5): stride-addressed CUDA-core
  kernels that read the fp32 batch as it is;
* anything else (grouped / dilated convolutions, fp32 compute with many channels) stays on ATen and is
  counted in ``STATS["aten"]``.

Backward: the input gradient stays on the calling stream (it is on the critical path of the backward
chain); the weight gradient runs on a side stream when a ``split_backward`` context is open, i.e. as a
parallel branch of the captured training graph, and lands in fp32 either in a per-layer sink of the
training step (``wgrad_sinks``; the fused SGD kernel reads it directly) or in an ordinary ``.grad``.
"""
import contextlib
import ctypes
import os
import threading
import types
from typing import Any, Iterator

import torch
from torch import nn

from . import runtime

_tls = threading.local()
STATS = {"tc": 0, "direct": 0, "stem": 0, "aten": 0}
_DTYPE = {torch.float32: 0, torch.bfloat16: 1}
_DIRECT_MAX_FILTER = 12288
_DIRECT_MAX_PATCH = 160  # Cin * R * S: stems and LeNet-5; wider layers need the tensor cores


class DlsConvShape(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int) for k in ("n", "h", "w", "c_in", "c_out", "r", "s", "stride", "pad_h", "pad_w",
                                            "x_pitch", "y_pitch", "dw_pitch", "dw_cols")]


class DlsDirectConv(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int) for k in ("n", "c_in", "h", "w", "c_out", "r", "s", "oh", "ow", "stride_h",
                                            "stride_w", "pad_h", "pad_w")] + \
               [("xs", ctypes.c_longlong * 4), ("ws", ctypes.c_longlong * 4), ("ys", ctypes.c_longlong * 4),
                ("x_dtype", ctypes.c_int), ("w_dtype", ctypes.c_int), ("y_dtype", ctypes.c_int)]


def _enabled() -> bool:
    return os.environ.get("DLS_OWN_CONV", "1") != "0"


def _is_nhwc(t: torch.Tensor) -> bool:
    """Physically [N][H][W][C] dense (a 1x1 spatial or 1-channel tensor qualifies in either format)."""
    n, c, h, w = t.shape
    want = (h * w * c, 1, w * c, c)
    return all(s == e for d, s, e in zip(t.shape, t.stride(), want) if d > 1)


def _nhwc(t: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    if t.dtype != dtype:
        t = t.to(dtype)
    if not _is_nhwc(t):
        t = t.contiguous(memory_format=torch.channels_last)
    return t


def _out_size(size: int, k: int, stride: int, pad: int) -> int:
    return (size + 2 * pad - k) // stride + 1


def tc_shape(x_shape: tuple, w_shape: tuple, stride: tuple, padding: tuple) -> DlsConvShape | None:
    """The tensor-core kernels' problem descriptor, or None when the shape is not eligible."""
    n, c, h, w = x_shape
    k, _, r, s = w_shape
    if stride[0] != stride[1] or stride[0] not in (1, 2):
        return None
    shape = DlsConvShape(n, h, w, c, k, r, s, stride[0], padding[0], padding[1], 0, 0, 0, 0)
    if stride[0] == 2 and (_out_size(h, r, 2, padding[0]) != h // 2 or _out_size(w, s, 2, padding[1]) != w // 2):
        return None
    if not runtime.require().dls_conv_supported(ctypes.byref(shape)):
        return None
    return shape


def _direct_desc(x: torch.Tensor, w: torch.Tensor, y: torch.Tensor, stride: tuple, padding: tuple) -> DlsDirectConv:
    n, c, h, wd = x.shape
    k, _, r, s = w.shape
    d = DlsDirectConv(n, c, h, wd, k, r, s, y.shape[2], y.shape[3], stride[0], stride[1], padding[0], padding[1])
    d.xs = (ctypes.c_longlong * 4)(*x.stride())
    d.ws = (ctypes.c_longlong * 4)(*w.stride())
    d.ys = (ctypes.c_longlong * 4)(*y.stride())
    d.x_dtype, d.w_dtype, d.y_dtype = _DTYPE[x.dtype], _DTYPE[w.dtype], _DTYPE[y.dtype]
    return d


def wgrad_needs_zero(shape: DlsConvShape) -> bool:
    """Whether ``dls_conv_wgrad`` would split the pixel range (and therefore accumulate with reductions)."""
    oh, ow = _out_size(shape.h, shape.r, shape.stride, shape.pad_h), _out_size(shape.w, shape.s, shape.stride, shape.pad_w)

    def p2(v: int) -> int:
        return 1 << max(v - 1, 0).bit_length()

    bw = min(p2(ow), 128)
    bh = min(p2(oh), 128 // bw)
    bn = 128 // (bw * bh)
    tiles = -(-ow // bw) * -(-oh // bh) * -(-shape.n // bn)
    return tiles >= 16


# ---- functional front-ends of the tensor-core kernels (NHWC bf16 tensors given as N,C,H,W views) -----------
def tc_fprop(x: torch.Tensor, weight: torch.Tensor, shape: DlsConvShape, bias: torch.Tensor | None = None,
             stats: torch.Tensor | None = None) -> torch.Tensor:
    """``conv(x, weight) + bias`` -> bf16 channels-last; ``stats`` (fp32 [2, Cout], zeroed by the caller)
    receives per-channel sum / sum of squares of the outputs."""
    oh, ow = _out_size(shape.h, shape.r, shape.stride, shape.pad_h), _out_size(shape.w, shape.s, shape.stride, shape.pad_w)
    y = torch.empty((shape.n, shape.c_out, oh, ow), dtype=torch.bfloat16, device=x.device,
                    memory_format=torch.channels_last)
    runtime.call("dls_conv_fprop", x.data_ptr(), weight.data_ptr(), y.data_ptr(), runtime.ptr(bias), runtime.ptr(stats),
                 ctypes.byref(shape), runtime.stream_ptr(x.device))
    return y


def tc_dgrad(dy: torch.Tensor, weight: torch.Tensor, shape: DlsConvShape) -> torch.Tensor:
    dx = torch.empty((shape.n, shape.c_in, shape.h, shape.w), dtype=torch.bfloat16, device=dy.device,
                     memory_format=torch.channels_last)
    runtime.call("dls_conv_dgrad", dy.data_ptr(), weight.data_ptr(), dx.data_ptr(), ctypes.byref(shape),
                 runtime.stream_ptr(dy.device))
    return dx


def tc_wgrad(dy: torch.Tensor, x: torch.Tensor, shape: DlsConvShape, out: torch.Tensor | None = None,
             accumulate: bool = True) -> torch.Tensor:
    """``[Cout, Cin, R, S]`` in channels-last physical order.  ``accumulate``: ``out`` (fp32) holds zeros or a
    running sum and the kernel may split the pixel range; otherwise every element is stored once, in the dtype of
    ``out`` (fp32 or bf16)."""
    if out is None:
        out = torch.zeros((shape.c_out, shape.r, shape.s, shape.c_in), dtype=torch.float32,
                          device=dy.device).permute(0, 3, 1, 2)
    assert out.dtype == torch.float32 or not accumulate
    mode = 1 if accumulate else (2 if out.dtype == torch.bfloat16 else 0)
    runtime.call("dls_conv_wgrad", dy.data_ptr(), x.data_ptr(), out.data_ptr(), ctypes.byref(shape), mode,
                 runtime.stream_ptr(dy.device))
    return out


def _stem_desc(x: torch.Tensor, w: torch.Tensor, oh: int, ow: int, stride: tuple, padding: tuple) -> DlsDirectConv:
    n, c, h, wd = x.shape
    k, _, r, s = w.shape
    d = DlsDirectConv(n, c, h, wd, k, r, s, oh, ow, stride[0], stride[1], padding[0], padding[1])
    d.xs = (ctypes.c_longlong * 4)(*x.stride())
    d.ws = (ctypes.c_longlong * 4)(*w.stride())
    d.x_dtype, d.w_dtype, d.y_dtype = _DTYPE[x.dtype], _DTYPE[w.dtype], 1
    return d


def take_stats(channels: int, device: torch.device) -> torch.Tensor:
    """A zeroed fp32 [2, channels] buffer for the fused BatchNorm statistics: the next slice of the training
    step's arena (zeroed once per step) when one is open, else a fresh tensor."""
    arena = getattr(_tls, "stats_arena", None)
    if arena is not None:
        buf, used = arena
        need = 2 * channels
        if used + need <= buf.numel():
            arena[1] = used + need
            return buf[used: used + need].view(2, channels)
    return torch.zeros(2, channels, dtype=torch.float32, device=device)


def take_arena_slice(floats: int) -> "torch.Tensor | None":
    """``floats`` zeroed fp32 values from the open per-step arena, or None when no arena is open / it is exhausted
    (callers then fall back to a path that needs no pre-zeroed memory)."""
    arena = getattr(_tls, "stats_arena", None)
    if arena is None:
        return None
    buf, used = arena
    if used + floats > buf.numel():
        return None
    arena[1] = used + floats
    return buf[used: used + floats]


class _Conv2d(torch.autograd.Function):
    """mode 1 = tensor cores, 2 = direct, 3 = stem (explicit patches + tensor cores).  ``sink``: persistent
    fp32 buffer in the filter's physical layout that receives the weight gradient instead of ``weight.grad``
    (training-step graphs).  Returns ``(y, stats)``; ``stats`` (or None) holds the per-channel sum / sum of
    squares of ``y`` for the BatchNorm that follows."""

    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, weight: torch.Tensor, bias: Any, stride: tuple, padding: tuple, mode: int,
                compute_dtype: torch.dtype, side: Any, sink: Any, sink_zeroed: bool, want_stats: bool) -> tuple:
        ctx.set_materialize_grads(False)  # no zero-filled gradient tensor for the statistics output
        ctx.conf = (stride, padding, mode, compute_dtype)
        ctx.side, ctx.sink, ctx.sink_zeroed = side, sink, sink_zeroed
        ctx.has_bias = bias is not None
        ctx.in_dtypes = (x.dtype, weight.dtype, None if bias is None else bias.dtype)
        ctx.x_meta = (tuple(x.shape), x.stride())
        n, c, h, wd = x.shape
        k, _, r, s = weight.shape
        oh, ow = _out_size(h, r, stride[0], padding[0]), _out_size(wd, s, stride[1], padding[1])
        stream = runtime.stream_ptr(x.device)
        stats = take_stats(k, x.device) if (want_stats and mode in (1, 3)) else None
        if mode == 1:
            xb = _nhwc(x, torch.bfloat16)
            wb = _nhwc(weight.detach(), torch.bfloat16)
            shape = tc_shape(tuple(x.shape), tuple(weight.shape), stride, padding)
            bias32 = None if bias is None else bias.detach().float().contiguous()
            y = tc_fprop(xb, wb, shape, bias32, stats)
            ctx.shape = shape
            ctx.save_for_backward(xb, wb)
            STATS["tc"] += 1
        elif mode == 3:
            q = c * r * s
            kp = (q + 7) // 8 * 8
            wd_ = weight.detach()
            patches = torch.empty((n * oh * ow, kp), dtype=torch.bfloat16, device=x.device)
            wpk = torch.empty((k, kp), dtype=torch.bfloat16, device=x.device)
            d = _stem_desc(x, wd_, oh, ow, stride, padding)
            runtime.call("dls_conv_stem_im2col", x.data_ptr(), wd_.data_ptr(), patches.data_ptr(), wpk.data_ptr(), kp,
                         ctypes.byref(d), stream)
            shape = DlsConvShape(n, oh, ow, kp, k, 1, 1, 1, 0, 0, 0, 0, q, q)
            bias32 = None if bias is None else bias.detach().float().contiguous()
            y = tc_fprop(patches, wpk, shape, bias32, stats)
            ctx.shape = shape
            ctx.save_for_backward(patches, wd_)
            STATS["stem"] += 1
        else:
            y = torch.empty((n, k, oh, ow), dtype=compute_dtype, device=x.device, memory_format=torch.channels_last)
            wd_ = weight.detach()
            bd = None if bias is None else bias.detach()
            d = _direct_desc(x, wd_, y, stride, padding)
            runtime.call("dls_conv_direct_fprop", x.data_ptr(), wd_.data_ptr(), runtime.ptr(bd),
                         0 if bd is None else _DTYPE[bd.dtype], y.data_ptr(), ctypes.byref(d), stream)
            ctx.save_for_backward(x, wd_)
            STATS["direct"] += 1
        if stats is not None:
            ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor, _dstats: Any = None) -> tuple:
        if dy is None:
            return (None,) * 11
        x, weight = ctx.saved_tensors  # mode 3: x holds the patch matrix
        stride, padding, mode, compute_dtype = ctx.conf
        side, sink = ctx.side, ctx.sink
        need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2]
        grad_x = grad_w = grad_b = None
        main = torch.cuda.current_stream(dy.device)
        if mode in (1, 3):
            dy = _nhwc(dy, torch.bfloat16)
        elif dy.dtype not in _DTYPE:
            dy = dy.float()

        def weight_gradient() -> tuple:
            stream = runtime.stream_ptr(dy.device)
            gb = None
            permute_back = False
            if sink is not None and (mode != 3 or _is_nhwc(weight)):
                buf, zeroed = sink, ctx.sink_zeroed
            elif mode == 3:
                # the stem kernel writes [Cout][R][S][Cin]
                k, c, r, s = weight.shape
                buf = torch.zeros((k, r, s, c), dtype=torch.float32, device=dy.device).permute(0, 3, 1, 2)
                zeroed, permute_back = True, not _is_nhwc(weight)
            else:
                # fp32 buffer with the physical layout of the filter
                buf = torch.zeros_like(weight, dtype=torch.float32)
                zeroed = True
            if mode in (1, 3):
                tc_wgrad(dy, x, ctx.shape, out=buf, accumulate=bool(zeroed))
                if need_b:
                    gb = dy.float().sum(dim=(0, 2, 3))
            else:
                assert zeroed, "the direct weight-gradient kernel accumulates"
                if need_b:
                    gb = torch.zeros(weight.shape[0], dtype=torch.float32, device=dy.device)
                d = _direct_desc(x, buf, dy, stride, padding)
                runtime.call("dls_conv_direct_wgrad", dy.data_ptr(), x.data_ptr(), buf.data_ptr(), runtime.ptr(gb),
                             ctypes.byref(d), stream)
            if buf is sink:
                gw = None
            else:
                gw = buf.to(ctx.in_dtypes[1])
                if permute_back:
                    gw = gw.contiguous()
            if gb is not None:
                gb = gb.to(ctx.in_dtypes[2])
            return gw, gb

        if need_w or need_b:
            if side is not None:
                side.wait_stream(main)  # fork: dy (and everything before it) is ready
                with torch.cuda.stream(side):
                    grad_w, grad_b = weight_gradient()
                dy.record_stream(side)
                x.record_stream(side)
            else:
                grad_w, grad_b = weight_gradient()
        if need_x:
            stream = runtime.stream_ptr(dy.device)
            if mode == 1:
                grad_x = tc_dgrad(dy, weight, ctx.shape)
            else:
                x_shape, x_stride = ctx.x_meta
                grad_x = torch.empty_strided(x_shape, x_stride, dtype=compute_dtype if mode == 3 else x.dtype,
                                             device=dy.device)
                d = _direct_desc(grad_x, weight, dy, stride, padding)
                runtime.call("dls_conv_direct_dgrad", dy.data_ptr(), weight.data_ptr(), grad_x.data_ptr(),
                             ctypes.byref(d), stream)
            if grad_x.dtype != ctx.in_dtypes[0]:
                grad_x = grad_x.to(ctx.in_dtypes[0])
        return grad_x, grad_w, grad_b, None, None, None, None, None, None, None, None


def _route(module: nn.Conv2d, x: torch.Tensor) -> tuple[int, torch.dtype]:
    """0 = ATen, 1 = tensor cores, 2 = direct, 3 = stem; and the compute dtype."""
    if not (_enabled() and x.is_cuda and x.dim() == 4 and runtime.available()):
        return 0, x.dtype
    if module.groups != 1 or module.dilation != (1, 1) or module.padding_mode != "zeros" \
            or isinstance(module.padding, str) or x.dtype not in _DTYPE or module.weight.dtype not in _DTYPE:
        return 0, x.dtype
    autocast = torch.is_autocast_enabled(x.device.type)
    dtype = torch.get_autocast_dtype(x.device.type) if autocast else x.dtype
    if autocast and dtype != torch.bfloat16:
        return 0, x.dtype
    w = module.weight
    patch = w.shape[1] * w.shape[2] * w.shape[3]
    if dtype == torch.bfloat16:
        if tc_shape(tuple(x.shape), tuple(w.shape), module.stride, module.padding) is not None:
            return 1, dtype
        if w.shape[1] < 8 and w.shape[0] % 8 == 0 and patch <= 256 and os.environ.get("DLS_CONV_STEM", "1") != "0":
            return 3, dtype
    if w.numel() <= _DIRECT_MAX_FILTER and patch <= _DIRECT_MAX_PATCH:
        return 2, dtype
    return 0, dtype


def forward_with_stats(self: nn.Conv2d, x: torch.Tensor, want_stats: bool = True) -> tuple:
    """``(conv(x), stats)``: stats is None unless the layer ran on the tensor-core kernels."""
    mode, dtype = _route(self, x)
    if mode == 0:
        STATS["aten"] += 1
        return self._conv_forward(x, self.weight, self.bias), None
    side = getattr(_tls, "side", None) if torch.is_grad_enabled() else None
    sinks = getattr(_tls, "sinks", None)
    sink = zeroed = None
    if sinks is not None and torch.is_grad_enabled():
        entry = sinks.get(id(self))
        if entry is not None:
            if entry[0] is None:
                # first forward of the layer inside a training step: decide whether its weight-gradient kernel
                # splits the pixel range (fp32 sink that accumulates) or stores once (sink in the filter's dtype)
                w = self.weight
                entry[1] = mode != 1 or wgrad_needs_zero(tc_shape(tuple(x.shape), tuple(w.shape), self.stride, self.padding))
                keep = torch.bfloat16 if (not entry[1] and w.dtype == torch.bfloat16) else torch.float32
                entry[0] = torch.zeros_like(w, dtype=keep)
            sink, zeroed = entry
    return _Conv2d.apply(x, self.weight, self.bias, tuple(self.stride), tuple(self.padding), mode, dtype, side, sink,
                         bool(zeroed), want_stats)


def _forward(self: nn.Conv2d, x: torch.Tensor) -> torch.Tensor:
    return forward_with_stats(self, x, want_stats=False)[0]


def install(model: nn.Module) -> int:
    """Route the ``nn.Conv2d`` layers of ``model`` through this framework's kernels.  Idempotent."""
    patched = 0
    for module in model.modules():
        if type(module) is nn.Conv2d:
            if not getattr(module, "_dls_own_conv", False):
                module.forward = types.MethodType(_forward, module)
                module._dls_own_conv = True
            patched += 1
    return patched


@contextlib.contextmanager
def split_backward(side: "torch.cuda.Stream | None", sinks: dict | None = None) -> Iterator[None]:
    """While open (forward pass of the calling thread), convolutions record ``side`` as the stream of their
    weight-gradient kernels and ``sinks`` ({id(module): [fp32 buffer, accumulate]}) as the destination.
    The caller joins: ``main.wait_stream(side)`` after backward."""
    previous = (getattr(_tls, "side", None), getattr(_tls, "sinks", None))
    _tls.side, _tls.sinks = side, sinks
    try:
        yield
    finally:
        _tls.side, _tls.sinks = previous


@contextlib.contextmanager
def stats_arena(buf: "torch.Tensor | None") -> Iterator[None]:
    """While open, ``take_stats`` hands out consecutive slices of ``buf`` (fp32, zeroed by the caller once per
    step) instead of allocating + zero-filling a tensor per layer."""
    previous = getattr(_tls, "stats_arena", None)
    _tls.stats_arena = None if buf is None else [buf, 0]
    try:
        yield
    finally:
        _tls.stats_arena = previous


def oracle_conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Any, stride: tuple, padding: tuple) -> torch.Tensor:
    """Plain PyTorch fp32 statement of the op (test oracle)."""
    return torch.nn.functional.conv2d(x.float(), weight.float(), None if bias is None else bias.float(), stride, padding)
