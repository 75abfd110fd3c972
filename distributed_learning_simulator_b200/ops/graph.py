"""Front-ends of csrc/graph.cu: mask-driven ordered compaction of edges and
nodes, boundary-embedding row gather, halo scatter-merge, and the normalised GCN aggregation.

Every function has a plain-PyTorch oracle (used on CPU and by the numerics tests); on a GPU box
the CUDA kernels are the path that runs.  Reference behaviour:
/root/reference/simulation_lib/worker/graph_worker.py:205-216 (per-batch edge masks), :252-269
(boundary rows), :271-321 (merge of received embeddings).
"""
from typing import Any

import torch

from . import runtime


def _use_kernels(t: torch.Tensor) -> bool:
    return t.is_cuda and runtime.available()


def _mask_u8(mask: torch.Tensor) -> torch.Tensor:
    assert mask.dtype in (torch.bool, torch.uint8)
    return mask.contiguous()


# ---- K13: ordered compaction ---------------------------------------------------------------------------
def oracle_keep(a: torch.Tensor, b: torch.Tensor | None, mask: torch.Tensor, n_id: torch.Tensor | None) -> torch.Tensor:
    ga = a if n_id is None else n_id[a]
    keep = mask[ga].bool()
    if b is not None:
        gb = b if n_id is None else n_id[b]
        keep = keep & mask[gb].bool()
    return keep


def compact_pairs(a: torch.Tensor, b: torch.Tensor | None, mask: torch.Tensor, n_id: torch.Tensor | None = None,
                  keep_prob: float = 1.0, seed: int = 0,
                  want_pos: bool = False) -> tuple[torch.Tensor, torch.Tensor | None, torch.Tensor | None]:
    """Stable selection of the pairs (a[i], b[i]) whose endpoints are set in ``mask`` (looked up
    through ``n_id`` when given), optionally thinned by a Bernoulli(keep_prob) draw per pair.
    Returns (a_kept, b_kept or None, original positions or None)."""
    n = a.numel()
    if not _use_kernels(a):
        assert keep_prob >= 1.0, "the CPU oracle has no Philox stream; thin the edges with torch.rand instead"
        keep = oracle_keep(a, b, mask, n_id)
        pos = keep.nonzero(as_tuple=False).view(-1)
        return a[keep], None if b is None else b[keep], pos if want_pos else None
    dev = a.device
    a = a.contiguous()
    b = None if b is None else b.contiguous()
    assert a.dtype == torch.int64 and (b is None or b.dtype == torch.int64)
    n_id = None if n_id is None else n_id.contiguous()
    mask = _mask_u8(mask)
    out_a = torch.empty(n, dtype=torch.int64, device=dev)
    out_b = None if b is None else torch.empty(n, dtype=torch.int64, device=dev)
    out_pos = torch.empty(n, dtype=torch.int64, device=dev) if want_pos else None
    count = torch.empty(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(n // 2048 + 2, dtype=torch.int32, device=dev)
    runtime.call("dls_index_compact", a.data_ptr(), runtime.ptr(b), n, runtime.ptr(n_id), mask.data_ptr(),
                 float(keep_prob), int(seed) & 0xFFFFFFFFFFFFFFFF, out_a.data_ptr(), runtime.ptr(out_b),
                 runtime.ptr(out_pos), count.data_ptr(), scratch.data_ptr(), runtime.stream_ptr(dev))
    k = int(count.item())  # the caller slices: same host read a boolean index would have cost
    return out_a[:k], None if out_b is None else out_b[:k], None if out_pos is None else out_pos[:k]


def filter_edges(edge_index: torch.Tensor, node_mask: torch.Tensor, n_id: torch.Tensor | None = None,
                 drop_rate: float = 0.0, seed: int = 0) -> torch.Tensor:
    """edge_index[:, keep] with keep = mask[src] & mask[dst] (& Bernoulli(1 - drop_rate))."""
    src, dst, _ = compact_pairs(edge_index[0], edge_index[1], node_mask, n_id, 1.0 - float(drop_rate), seed)
    return torch.stack([src, dst])


def select_nodes(n_id: torch.Tensor, node_mask: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """(rows, n_id[rows]) for the rows of a mini-batch whose global id is set in ``node_mask``."""
    ids, _, rows = compact_pairs(n_id, None, node_mask, None, want_pos=True)
    return rows, ids


# ---- K14a: row gather -------------------------------------------------------------------------------------
def gather_rows(x: torch.Tensor, rows: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """x.index_select(0, rows) (detached); ``out`` may live on a peer GPU (P2P-mapped)."""
    x = x.detach()
    if not _use_kernels(x) or (x.shape[1] * x.element_size()) % 4:
        res = x.index_select(0, rows)
        if out is not None:
            out.copy_(res)
            return out
        return res
    x = x.contiguous()
    if out is None:
        out = torch.empty((rows.numel(), x.shape[1]), dtype=x.dtype, device=x.device)
    assert out.is_contiguous() and out.shape == (rows.numel(), x.shape[1]) and out.dtype == x.dtype
    runtime.call("dls_row_gather", x.data_ptr(), rows.contiguous().data_ptr(), rows.numel(),
                 x.shape[1] * x.element_size(), out.data_ptr(), runtime.stream_ptr(x.device))
    return out


# ---- K14b: halo merge ---------------------------------------------------------------------------------------
def oracle_halo_merge(x: torch.Tensor, n_id: torch.Tensor, train_mask: torch.Tensor, ids: torch.Tensor | None,
                      emb: torch.Tensor | None) -> torch.Tensor:
    foreign = ~train_mask.bool()[n_id]
    new_x = torch.where(foreign.unsqueeze(-1), torch.zeros_like(x), x)
    if emb is None or ids is None or ids.numel() == 0:
        return new_x
    order = torch.argsort(ids)
    pos = torch.searchsorted(ids[order], n_id).clamp(max=ids.numel() - 1)
    hit = (ids[order][pos] == n_id) & foreign
    rows = hit.nonzero(as_tuple=False).view(-1)
    return new_x.index_copy(0, rows, emb[order[pos[rows]]].to(x.dtype))


class _HaloMerge(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, n_id: torch.Tensor, train_mask: torch.Tensor, sorted_ids: Any,
                perm: Any, emb: Any) -> torch.Tensor:
        x = x.contiguous()
        out = torch.empty_like(x)
        foreign = torch.empty(x.shape[0], dtype=torch.uint8, device=x.device)
        n_ids = 0 if sorted_ids is None else sorted_ids.numel()
        runtime.call("dls_halo_merge", x.data_ptr(), n_id.data_ptr(), x.shape[0], x.shape[1] * x.element_size(),
                     train_mask.data_ptr(), runtime.ptr(sorted_ids), runtime.ptr(perm), n_ids, runtime.ptr(emb),
                     None, foreign.data_ptr(), out.data_ptr(), runtime.stream_ptr(x.device))
        ctx.save_for_backward(foreign)
        return out

    @staticmethod
    def backward(ctx: Any, grad: torch.Tensor) -> tuple:
        (foreign,) = ctx.saved_tensors
        grad = grad.contiguous()
        out = torch.empty_like(grad)
        runtime.call("dls_halo_merge", grad.data_ptr(), None, grad.shape[0], grad.shape[1] * grad.element_size(),
                     None, None, None, 0, None, foreign.data_ptr(), None, out.data_ptr(),
                     runtime.stream_ptr(grad.device))
        return out, None, None, None, None, None


def halo_merge(x: torch.Tensor, n_id: torch.Tensor, train_mask: torch.Tensor, ids: torch.Tensor | None,
               emb: torch.Tensor | None) -> torch.Tensor:
    """Rows of foreign nodes are replaced by the received embedding of that node (or zero); rows
    of local training nodes pass through with their gradient."""
    if not _use_kernels(x) or (x.shape[1] * x.element_size()) % 4:
        return oracle_halo_merge(x, n_id, train_mask, ids, emb)
    sorted_ids = perm = None
    if emb is not None and ids is not None and ids.numel() > 0:
        sorted_ids, perm = torch.sort(ids.to(x.device))
        emb = emb.to(device=x.device, dtype=x.dtype, non_blocking=True).contiguous()
    else:
        emb = None
    return _HaloMerge.apply(x, n_id.contiguous(), _mask_u8(train_mask), sorted_ids, perm, emb)


# ---- GCN aggregation --------------------------------------------------------------------------------------------
def oracle_gcn_aggregate(h: torch.Tensor, edge_index: torch.Tensor, bias: torch.Tensor | None) -> torch.Tensor:
    n = h.shape[0]
    loop = torch.arange(n, device=edge_index.device)
    ei = torch.cat([edge_index, torch.stack([loop, loop])], dim=1)
    deg = torch.zeros(n, dtype=h.dtype, device=h.device).index_add_(
        0, ei[1], torch.ones(ei.shape[1], dtype=h.dtype, device=h.device))
    dinv = deg.clamp(min=1.0).rsqrt()
    w = dinv[ei[0]] * dinv[ei[1]]
    out = torch.zeros_like(h).index_add_(0, ei[1], h[ei[0]] * w.unsqueeze(-1))
    return out if bias is None else out + bias


class _GCNAggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, h: torch.Tensor, edge_index: torch.Tensor, bias: Any) -> torch.Tensor:
        h = h.contiguous()
        n, c = h.shape
        dev = h.device
        src, dst = edge_index[0].contiguous(), edge_index[1].contiguous()
        deg = torch.empty(n, dtype=torch.int32, device=dev)
        dinv = torch.empty(n, dtype=torch.float32, device=dev)
        stream = runtime.stream_ptr(dev)
        runtime.call("dls_gcn_dinv", dst.data_ptr(), dst.numel(), n, deg.data_ptr(), dinv.data_ptr(), stream)
        out = torch.empty_like(h)
        runtime.call("dls_gcn_aggregate", h.data_ptr(), src.data_ptr(), dst.data_ptr(), src.numel(), n, c,
                     dinv.data_ptr(), runtime.ptr(bias), out.data_ptr(), stream)
        ctx.save_for_backward(src, dst, dinv)
        ctx.has_bias = bias is not None
        return out

    @staticmethod
    def backward(ctx: Any, grad: torch.Tensor) -> tuple:
        src, dst, dinv = ctx.saved_tensors
        grad = grad.contiguous()
        n, c = grad.shape
        # transposed operator: messages flow dst -> src with the same symmetric weights
        grad_h = torch.empty_like(grad)
        runtime.call("dls_gcn_aggregate", grad.data_ptr(), dst.data_ptr(), src.data_ptr(), src.numel(), n, c,
                     dinv.data_ptr(), None, grad_h.data_ptr(), runtime.stream_ptr(grad.device))
        return grad_h, None, grad.sum(dim=0) if ctx.has_bias else None


def gcn_aggregate(h: torch.Tensor, edge_index: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """D^-1/2 (A + I) D^-1/2 h (+ bias) for a COO ``edge_index`` (messages flow row 0 -> row 1)."""
    if not _use_kernels(h) or h.dtype != torch.float32:
        return oracle_gcn_aggregate(h, edge_index, bias)
    return _GCNAggregate.apply(h, edge_index, None if bias is None else bias.contiguous())
