"""Front-ends of csrc/misc.cu: sign packing / majority vote (sign_SGD), Bernoulli upload
masking (FedDropoutAvg), per-block squared parameter change (FedOBD)."""
from typing import Any

import torch

from . import runtime

_seed = [0x243F6A8885A308D3]


def _next_seed() -> int:
    _seed[0] = (_seed[0] * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
    return _seed[0]


def sign_pack(grad: torch.Tensor) -> torch.Tensor:
    n = grad.numel()
    if grad.is_cuda and runtime.available():
        packed = torch.empty((n + 31) // 32, dtype=torch.int32, device=grad.device)
        runtime.call("dls_sign_pack", grad.data_ptr(), n, packed.data_ptr(), runtime.stream_ptr(grad.device))
        return packed
    return oracle_sign_pack(grad)


def oracle_sign_pack(grad: torch.Tensor) -> torch.Tensor:
    n = grad.numel()
    bits = (grad.reshape(-1) < 0).to(torch.int64)
    pad = (-n) % 32
    if pad:
        bits = torch.cat([bits, bits.new_zeros(pad)])
    weights = (1 << torch.arange(32, dtype=torch.int64, device=grad.device))
    words = (bits.view(-1, 32) * weights).sum(dim=1)
    words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)
    return words.to(torch.int32)


def sign_majority(rows: list[torch.Tensor], n: int) -> torch.Tensor:
    device = rows[0].device
    if rows[0].is_cuda and runtime.available():
        rows = [r.to(device, non_blocking=True) for r in rows]
        table = runtime.pointer_table(rows, device)
        vote = torch.empty(n, dtype=torch.float32, device=device)
        runtime.call("dls_sign_majority", table.data_ptr(), len(rows), n, vote.data_ptr(), runtime.stream_ptr(device))
        return vote
    return oracle_sign_majority(rows, n)


class PeerWords:
    """Packed sign words that live in another rank's peer-mapped memory: usable by this framework's kernels
    through ``data_ptr()``; ``to_local()`` copies them here for code that needs a tensor."""

    def __init__(self, ptr: int, n_words: int, device: torch.device) -> None:
        self.ptr, self.n_words, self.device = int(ptr), int(n_words), device

    def data_ptr(self) -> int:
        return self.ptr

    def to_local(self) -> torch.Tensor:
        out = torch.empty(self.n_words, dtype=torch.int32, device=self.device)
        runtime.call("dls_copy_device", out.data_ptr(), self.ptr, self.n_words * 4, runtime.stream_ptr(self.device))
        return out


def as_local_words(row: Any) -> torch.Tensor:
    return row.to_local() if isinstance(row, PeerWords) else row


def sign_vote_sgd(rows: list, n: int, params: torch.Tensor, mom: torch.Tensor | None, lr: float, momentum: float,
                  dampening: float = 0.0, nesterov: bool = False, mom_ready: bool = True,
                  shadow: torch.Tensor | None = None) -> None:
    """One launch: majority vote over the packed rows (tensors or raw device pointers, e.g. peer-mapped words of
    other ranks) and the SGD update of ``params`` (flat fp32) with it.  CPU / no library: the torch oracle."""
    if params.is_cuda and runtime.available():
        table = runtime.pointer_table(rows, params.device)
        runtime.call("dls_sign_vote_sgd", table.data_ptr(), len(rows), int(n), params.data_ptr(), runtime.ptr(mom),
                     float(lr), float(momentum), float(dampening), int(nesterov), int(mom_ready), runtime.ptr(shadow),
                     None, None, 0, runtime.stream_ptr(params.device))
        return
    d = oracle_sign_majority([r for r in rows], n)
    if momentum != 0:
        assert mom is not None
        if mom_ready:
            mom.mul_(momentum).add_(d, alpha=1 - dampening)
        else:
            mom.copy_(d)
        d = d + momentum * mom if nesterov else mom
    params.add_(d, alpha=-lr)
    if shadow is not None:
        shadow.copy_(params)


def oracle_sign_majority(rows: list[torch.Tensor], n: int) -> torch.Tensor:
    total = torch.zeros(n, dtype=torch.int64, device=rows[0].device)
    shifts = torch.arange(32, dtype=torch.int64, device=rows[0].device)
    for r in rows:
        w = r.to(torch.int64) & 0xFFFFFFFF
        neg = ((w.view(-1, 1) >> shifts) & 1).view(-1)[:n]
        total += 1 - 2 * neg
    return torch.sign(total).float()


def dropout_mask(flat: torch.Tensor, keep_prob: float, seed: int | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns (masked copy, device scalar with the number of non-zero elements)."""
    if flat.is_cuda and runtime.available():
        out = torch.empty_like(flat)
        count = torch.zeros(1, dtype=torch.int64, device=flat.device)
        runtime.call("dls_dropout_mask", flat.data_ptr(), flat.numel(), float(keep_prob),
                     _next_seed() if seed is None else int(seed), out.data_ptr(), count.data_ptr(),
                     runtime.stream_ptr(flat.device))
        return out, count
    mask = torch.bernoulli(torch.full_like(flat, keep_prob))
    out = flat * mask
    return out, torch.count_nonzero(out).reshape(1)


def block_sqdiff(cur: torch.Tensor, prev: torch.Tensor, tensor_offsets: torch.Tensor, block_of_tensor: torch.Tensor,
                 n_blocks: int) -> torch.Tensor:
    """scores[b] = sum over the tensors of block b of (cur - prev)^2 (fp64)."""
    if cur.is_cuda and runtime.available():
        scores = torch.empty(n_blocks, dtype=torch.float64, device=cur.device)
        runtime.call("dls_block_sqdiff", cur.data_ptr(), prev.data_ptr(), tensor_offsets.data_ptr(),
                     block_of_tensor.data_ptr(), block_of_tensor.numel(), n_blocks, scores.data_ptr(),
                     runtime.stream_ptr(cur.device))
        return scores
    return oracle_block_sqdiff(cur, prev, tensor_offsets, block_of_tensor, n_blocks)


def oracle_block_sqdiff(cur, prev, tensor_offsets, block_of_tensor, n_blocks) -> torch.Tensor:
    scores = torch.zeros(n_blocks, dtype=torch.float64, device=cur.device)
    offs = tensor_offsets.tolist()
    blocks = block_of_tensor.tolist()
    d = (cur - prev).double() ** 2  # fp32 difference, fp64 accumulation (what the kernel does)
    for t, b in enumerate(blocks):
        if b >= 0:
            scores[b] += d[offs[t]: offs[t + 1]].sum()
    return scores


def im2col_grouped(x_nhwc: torch.Tensor, groups: int, k: int, channels_last_k: bool, k_pad: int | None = None,
                   pad: int = 0) -> torch.Tensor:
    """x_nhwc: [B, H, W, groups * C] contiguous.  Returns cols [groups, B*OH*OW, k_pad] for k x k windows with
    ``pad`` pixels of zero padding; column order (kh, kw, c) if ``channels_last_k`` else (c, kh, kw)."""
    b, h, w, gc = x_nhwc.shape
    c = gc // groups
    assert gc == groups * c and x_nhwc.is_contiguous()
    kk = c * k * k
    if k_pad is None:
        k_pad = kk + (-kk) % 8
    oh, ow = h + 2 * pad - k + 1, w + 2 * pad - k + 1
    if x_nhwc.is_cuda and runtime.available() and x_nhwc.element_size() in (2, 4):
        cols = torch.empty((groups, b * oh * ow, k_pad), dtype=x_nhwc.dtype, device=x_nhwc.device)
        if pad == 0:
            runtime.call("dls_im2col_grouped", x_nhwc.data_ptr(), x_nhwc.element_size(), b, h, w, groups, c, k, k_pad,
                         int(channels_last_k), cols.data_ptr(), runtime.stream_ptr(x_nhwc.device))
        else:
            runtime.call("dls_im2col_grouped_padded", x_nhwc.data_ptr(), x_nhwc.element_size(), b, h, w, groups, c, k,
                         k_pad, int(channels_last_k), int(pad), cols.data_ptr(), runtime.stream_ptr(x_nhwc.device))
        return cols
    return oracle_im2col_grouped(x_nhwc, groups, k, channels_last_k, k_pad, pad)


def oracle_im2col_grouped(x_nhwc: torch.Tensor, groups: int, k: int, channels_last_k: bool, k_pad: int,
                          pad: int = 0) -> torch.Tensor:
    b, h, w, gc = x_nhwc.shape
    c = gc // groups
    oh, ow = h + 2 * pad - k + 1, w + 2 * pad - k + 1
    cols = torch.nn.functional.unfold(x_nhwc.permute(0, 3, 1, 2).float(), k, padding=pad).view(b, groups, c, k * k, oh * ow)
    cols = cols.permute(1, 0, 4, 3, 2) if channels_last_k else cols.permute(1, 0, 4, 2, 3)
    cols = cols.reshape(groups, b * oh * ow, c * k * k).to(x_nhwc.dtype)
    return torch.nn.functional.pad(cols, (0, k_pad - c * k * k))
