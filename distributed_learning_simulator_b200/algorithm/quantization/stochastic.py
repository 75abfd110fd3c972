"""QSGD-style unbiased stochastic quantization (FedPAQ, fed_obd_sq).

The reference imports ``stochastic_quantization(quantization_level=255)`` from the
external ``cyy_torch_algorithm`` package
(/root/reference/simulation_lib/topology/quantized_endpoint.py:8,76,82); its body is
not in the reference tree, so the oracle is defined here:
per tensor ``s = max|x|``, ``y = |x|/s * L``, ``q = floor(y) + Bernoulli(y - floor(y))``
stored as one byte, plus a sign bit plane; ``dequant = sign * q * s / L``; ``E[dequant] = x``.
The max-norm (rather than the L2 norm) is used because with L = 255 levels it keeps the
per-element error below ``s/255`` independent of tensor size.

On a CUDA flat payload the work is done by the hand-written ``sq_quant`` kernel
(csrc/quant.cu: segmented abs-max, Philox stochastic rounding, byte + sign-bit packing
in one pass); everywhere else by the torch code below, which is also the test oracle.
"""
from typing import Any, Callable

import torch

from ...engine.model_util import ParameterLayout
from ...engine.tensor import TensorDict
from ...message import FlatParameter, as_flat
from .packed import QuantizedFlat, pack_bits, unpack_bits


def sq_quantize_flat(flat: torch.Tensor, layout: ParameterLayout, level: int = 255,
                     generator: torch.Generator | None = None, seed: int | None = None) -> QuantizedFlat:
    if flat.is_cuda:
        from ...ops import quant as quant_ops

        if quant_ops.available():
            return quant_ops.sq_quant(flat, layout, level, seed)
    t = len(layout.names)
    offs = layout.offsets
    payload = torch.empty(layout.total, dtype=torch.uint8, device=flat.device)
    scale = torch.zeros(t, dtype=torch.float32, device=flat.device)
    for i in range(t):
        x = flat[offs[i] : offs[i + 1]].float()
        s = x.abs().max() if x.numel() else x.new_zeros(())
        scale[i] = s / level
        y = torch.where(s > 0, x.abs() / s.clamp(min=1e-38) * level, torch.zeros_like(x))
        lo = y.floor()
        r = torch.rand(y.shape, device=y.device, generator=generator)
        q = lo + (r < (y - lo)).float()
        payload[offs[i] : offs[i + 1]] = q.clamp(max=level).to(torch.uint8)
    # one sign-bit plane over the whole flat vector: byte j covers elements [8j, 8j+8)
    aux = pack_bits((flat < 0).to(torch.int64), 1)
    return QuantizedFlat(kind="sq", layout=layout, payload=payload, aux=aux, scale=scale,
                         offset=torch.zeros_like(scale), bits=torch.full((t,), 8, dtype=torch.int32, device=flat.device),
                         byte_offsets=torch.tensor(offs, dtype=torch.int64, device=flat.device))


def sq_dequantize_flat(q: QuantizedFlat, out: torch.Tensor | None = None) -> torch.Tensor:
    if q.payload.is_cuda:
        from ...ops import quant as quant_ops

        if quant_ops.available():
            return quant_ops.sq_dequant(q, out)
    layout = q.layout
    if out is None:
        out = torch.empty(layout.total, dtype=torch.float32, device=q.payload.device)
    neg = unpack_bits(q.aux, 1, layout.total).bool()
    for i in range(len(layout.names)):
        a, b = layout.offsets[i], layout.offsets[i + 1]
        mag = q.payload[a:b].float() * q.scale[i]
        out[a:b] = torch.where(neg[a:b], -mag, mag)
    return out


def stochastic_quantization(quantization_level: int = 255) -> tuple[Callable, Callable]:
    """Returns ``(quant, dequant)`` over TensorDicts (the reference's call shape)."""

    def quant(parameter: TensorDict) -> Any:
        fp = as_flat(parameter)
        if fp is not None:
            flat, layout = fp.flat, fp.layout
        else:
            layout = ParameterLayout.from_dict(parameter)
            flat = layout.flatten(parameter)
        return {"__quantized__": sq_quantize_flat(flat, layout, quantization_level)}

    def dequant(data: Any) -> Any:
        if isinstance(data, dict) and "__quantized__" in data:
            q = data["__quantized__"]
            return FlatParameter(sq_dequantize_flat(q), q.layout)
        return data

    return quant, dequant
