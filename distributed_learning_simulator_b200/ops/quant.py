"""Front-ends of the quantizer kernels (csrc/quant.cu).  Oracles live in
``algorithm/quantization`` (torch)."""
import torch

from ..engine.model_util import ParameterLayout
from . import runtime
from .reduce import layout_tables

_seed_counter = [0x9E3779B97F4A7C15]


def available() -> bool:
    return runtime.available()


def _next_seed() -> int:
    _seed_counter[0] = (_seed_counter[0] * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
    return _seed_counter[0]


def sq_quant(flat: torch.Tensor, layout: ParameterLayout, level: int = 255, seed: int | None = None):
    from ..algorithm.quantization.packed import QuantizedFlat

    device = flat.device
    n = layout.total
    t = len(layout.names)
    t_off, _ = layout_tables(layout, device)
    payload = torch.empty(n, dtype=torch.uint8, device=device)
    sign = torch.empty((n + 7) // 8, dtype=torch.uint8, device=device)
    scale = torch.empty(t, dtype=torch.float32, device=device)
    scratch = torch.empty(t, dtype=torch.int32, device=device)
    runtime.call("dls_sq_quant", flat.data_ptr(), n, t_off.data_ptr(), t, int(level),
                 _next_seed() if seed is None else int(seed), payload.data_ptr(), sign.data_ptr(), scale.data_ptr(),
                 scratch.data_ptr(), runtime.stream_ptr(device))
    return QuantizedFlat(kind="sq", layout=layout, payload=payload, aux=sign, scale=scale,
                         offset=torch.zeros_like(scale), bits=torch.full((t,), 8, dtype=torch.int32, device=device),
                         byte_offsets=t_off)


def sq_dequant(q, out: torch.Tensor | None = None) -> torch.Tensor:
    device = q.payload.device
    layout = q.layout
    if out is None:
        out = torch.empty(layout.total, dtype=torch.float32, device=device)
    t_off, _ = layout_tables(layout, device)
    runtime.call("dls_sq_dequant", q.payload.data_ptr(), q.aux.data_ptr(), q.scale.data_ptr(), t_off.data_ptr(),
                 len(layout.names), layout.total, out.data_ptr(), runtime.stream_ptr(device))
    return out


def nnadq_quant(flat: torch.Tensor, layout: ParameterLayout, weight: float):
    from ..algorithm.quantization.packed import QuantizedFlat

    device = flat.device
    t = len(layout.names)
    t_off, g_off = layout_tables(layout, device)
    groups = sum((n + 7) // 8 for n in layout.numels)
    capacity = groups * 16  # worst case: 16 bits everywhere
    payload = torch.empty(capacity, dtype=torch.uint8, device=device)
    scale = torch.empty(t, dtype=torch.float32, device=device)
    offset = torch.empty(t, dtype=torch.float32, device=device)
    bits = torch.empty(t, dtype=torch.int32, device=device)
    byte_offsets = torch.empty(t + 1, dtype=torch.int64, device=device)
    scratch = torch.empty(t * 4, dtype=torch.float64, device=device)
    runtime.call("dls_nnadq_quant", flat.data_ptr(), t_off.data_ptr(), g_off.data_ptr(), t, float(weight),
                 scratch.data_ptr(), scale.data_ptr(), offset.data_ptr(), bits.data_ptr(), byte_offsets.data_ptr(),
                 payload.data_ptr(), capacity, runtime.stream_ptr(device))
    q = QuantizedFlat(kind="nnadq", layout=layout, payload=payload, aux=torch.empty(0, dtype=torch.uint8, device=device),
                      scale=scale, offset=offset, bits=bits, byte_offsets=byte_offsets)
    q.lazy_payload_bytes = True  # real size is byte_offsets[-1]; read lazily (one sync) when asked
    return q


def nnadq_dequant(q, out: torch.Tensor | None = None) -> torch.Tensor:
    device = q.payload.device
    layout = q.layout
    if out is None:
        out = torch.empty(layout.total, dtype=torch.float32, device=device)
    t_off, g_off = layout_tables(layout, device)
    runtime.call("dls_nnadq_dequant", q.payload.data_ptr(), q.scale.data_ptr(), q.offset.data_ptr(),
                 q.bits.data_ptr(), q.byte_offsets.data_ptr(), t_off.data_ptr(), g_off.data_ptr(),
                 len(layout.names), out.data_ptr(), runtime.stream_ptr(device))
    return out
