"""Convolution whose backward runs its two halves on different streams.

Inside the captured training step every kernel depends on the previous one, although the weight
gradient of a convolution is needed only by the optimizer at the very end while the input gradient
is on the critical path of the backward chain.  ``split_backward(model, side_stream)`` makes every
``nn.Conv2d`` of ``model`` compute ``grad_weight`` on ``side_stream`` (forked from, and later joined
to, the main stream), so the CUDA graph gets a parallel branch: the small wgrad kernels (64 x 64
tiles, a fraction of the SMs) overlap the dgrad / BatchNorm-backward chain.  Same ATen/cuDNN kernels,
same numbers; only the dependency structure changes.  Used by ``engine/graph_step.py`` at capture
time (graph replays never run Python, so nothing is patched at run time).

The reference has no counterpart: its training loop is the external eager ``Trainer.train()``
(/root/reference/simulation_lib/worker/worker.py:86-88).
"""
import contextlib
import threading
from typing import Any, Iterator

import torch
from torch import nn

_tls = threading.local()


class _SplitBackwardConv2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, weight: torch.Tensor, bias: Any, stride: tuple, padding: tuple,
                dilation: tuple, groups: int, side: torch.cuda.Stream) -> torch.Tensor:
        ctx.in_dtypes = (x.dtype, weight.dtype, None if bias is None else bias.dtype)
        if torch.is_autocast_enabled(x.device.type):
            # what autocast would do to conv2d's arguments; the casts are no-ops for bf16 activations and
            # for weights that already live in the bf16 shadow arena
            dtype = torch.get_autocast_dtype(x.device.type)
            x, weight = x.to(dtype), weight.to(dtype)
            bias = None if bias is None else bias.to(dtype)
        ctx.save_for_backward(x, weight)
        ctx.conf = (stride, padding, dilation, groups)
        ctx.side = side
        ctx.has_bias = bias is not None
        with torch.autocast(x.device.type, enabled=False):
            return torch.nn.functional.conv2d(x, weight, bias, stride, padding, dilation, groups)

    @staticmethod
    def backward(ctx: Any, grad_out: torch.Tensor) -> tuple:
        x, weight = ctx.saved_tensors
        stride, padding, dilation, groups = ctx.conf
        side = ctx.side
        grad_x = grad_w = grad_b = None
        main = torch.cuda.current_stream(grad_out.device)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            side.wait_stream(main)  # fork: grad_out (and everything before it) is ready
            with torch.cuda.stream(side):
                _, grad_w, grad_b = torch.ops.aten.convolution_backward(
                    grad_out, x, weight, [weight.shape[0]] if ctx.has_bias else None, stride, padding, dilation,
                    False, [0, 0], groups, [False, bool(ctx.needs_input_grad[1]),
                                            bool(ctx.has_bias and ctx.needs_input_grad[2])])
                if grad_w is not None:
                    grad_w = grad_w.to(ctx.in_dtypes[1])
                if grad_b is not None:
                    grad_b = grad_b.to(ctx.in_dtypes[2])
            # the side stream reads tensors owned by the main stream's allocator pool
            grad_out.record_stream(side)
            x.record_stream(side)
            weight.record_stream(side)
        if ctx.needs_input_grad[0]:
            grad_x = torch.ops.aten.convolution_backward(
                grad_out, x, weight, None, stride, padding, dilation, False, [0, 0], groups, [True, False, False])[0]
            grad_x = grad_x.to(ctx.in_dtypes[0])
        return grad_x, grad_w, grad_b, None, None, None, None, None


def _forward(self: nn.Conv2d, x: torch.Tensor) -> torch.Tensor:
    side = getattr(_tls, "side", None)
    if side is None or not x.is_cuda or not torch.is_grad_enabled() or self.padding_mode != "zeros" \
            or isinstance(self.padding, str):
        return self._conv_forward(x, self.weight, self.bias)
    return _SplitBackwardConv2d.apply(x, self.weight, self.bias, self.stride, self.padding, self.dilation,
                                      self.groups, side)


def install(model: nn.Module) -> int:
    """Route the ``nn.Conv2d`` layers of ``model`` through the split-backward function (inactive
    until a ``split_backward`` context is open on the calling thread)."""
    import types

    patched = 0
    for module in model.modules():
        if type(module) is nn.Conv2d and not getattr(module, "_dls_split_backward", False):
            module.forward = types.MethodType(_forward, module)
            module._dls_split_backward = True
            patched += 1
    return patched


@contextlib.contextmanager
def split_backward(side: torch.cuda.Stream | None) -> Iterator[None]:
    """While open (forward pass of the calling thread), patched convolutions record ``side`` as the
    stream of their weight-gradient kernels.  The caller joins: ``main.wait_stream(side)`` after backward."""
    previous = getattr(_tls, "side", None)
    _tls.side = side
    try:
        yield
    finally:
        _tls.side = previous
