"""Packed representation shared by both quantizers.

A quantized model is ONE object over the flat parameter vector: a packed byte stream,
per-tensor metadata (scale/offset/bit-width), and the layout.  This is what travels in a
message and what the fused aggregation kernel decodes at load time (SURVEY.md K3/K4:
"dequant fused into K1's load path").  Bit order (identical in the CUDA kernels and the
torch oracle): element ``i`` of a tensor owns bits ``[i*b, (i+1)*b)`` of that tensor's
little-endian bit stream; every tensor's stream starts on a byte boundary and is padded
to a multiple of ``b`` bytes (8 elements).
"""
from dataclasses import dataclass

import torch

from ...engine.model_util import ParameterLayout


@dataclass
class QuantizedFlat:
    kind: str                    # "sq" (stochastic 255-level) | "nnadq"
    layout: ParameterLayout
    payload: torch.Tensor        # uint8 packed magnitudes / codes
    aux: torch.Tensor            # sq: packed sign bits (uint8); nnadq: empty
    scale: torch.Tensor          # float32 [T]  (sq: norm/level ; nnadq: step)
    offset: torch.Tensor         # float32 [T]  (nnadq: min ; sq: zeros)
    bits: torch.Tensor           # int32 [T]    (sq: 8)
    byte_offsets: torch.Tensor   # int64 [T+1]  start of each tensor in ``payload``

    @property
    def device(self) -> torch.device:
        return self.payload.device

    def compressed_bytes(self) -> int:
        t = len(self.layout.names)
        payload_bytes = self.payload.numel()
        if getattr(self, "lazy_payload_bytes", False):
            payload_bytes = int(self.byte_offsets[-1].item())  # CUDA path over-allocates the buffer
        return int(payload_bytes + self.aux.numel() + 4 * 3 * t)

    def original_bytes(self) -> int:
        return 4 * self.layout.total

    def compression_ratio(self) -> float:
        return self.compressed_bytes() / max(self.original_bytes(), 1)


def pack_bits(codes: torch.Tensor, bits: int) -> torch.Tensor:
    """Oracle bit packer: ``codes`` (int64 ``[n]``, values < 2**bits) -> uint8 ``[ceil(n/8)*bits]``."""
    n = codes.numel()
    pad = (-n) % 8
    if pad:
        codes = torch.cat([codes, codes.new_zeros(pad)])
    shifts = torch.arange(bits, device=codes.device, dtype=torch.int64)
    bit = ((codes.view(-1, 1) >> shifts) & 1).to(torch.uint8).view(-1)      # stream, LSB first
    bit = bit.view(-1, 8)
    weights = (1 << torch.arange(8, device=codes.device, dtype=torch.int64)).to(torch.uint8)
    return (bit * weights).sum(dim=1, dtype=torch.int64).to(torch.uint8)


def unpack_bits(packed: torch.Tensor, bits: int, n: int) -> torch.Tensor:
    shifts = torch.arange(8, device=packed.device, dtype=torch.int64)
    bit = ((packed.to(torch.int64).view(-1, 1) >> shifts) & 1).view(-1)
    total = ((n + 7) // 8) * 8
    bit = bit[: total * bits].view(total, bits)
    weights = 1 << torch.arange(bits, device=packed.device, dtype=torch.int64)
    return (bit * weights).sum(dim=1)[:n]
