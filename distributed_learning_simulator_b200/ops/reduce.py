"""Front-end of the fused FedAvg reduction kernels (csrc/reduce.cu).

``fused_aggregate`` turns a round's uploads -- flat fp32 rows, stochastic-quantized or
NNADQ-packed rows, full parameters or deltas, optionally with dropped blocks -- into the
new global model with ONE kernel launch (reference path: per-tensor Python loops on fp64
CPU tensors, simulation_lib/algorithm/fed_avg_algorithm.py:39-81).  ``oracle_aggregate``
is the plain-PyTorch statement of the same maths used by the tests.
"""
import ctypes
from typing import Any

import torch

from ..engine.model_util import FlatParameter, ParameterLayout, as_flat
from . import runtime

_group_cache: dict[tuple, tuple[torch.Tensor, torch.Tensor]] = {}
# how many aggregations each fused path has served in this process (tests and bench read it)
STATS = {"f32": 0, "packed": 0, "p2p": 0}


def layout_tables(layout: ParameterLayout, device: torch.device) -> tuple[torch.Tensor, torch.Tensor]:
    """(tensor_offsets[T+1], group_offsets[T+1]) on ``device`` (8-element groups per tensor)."""
    key = (hash(layout), layout.total, str(device))
    hit = _group_cache.get(key)
    if hit is None:
        groups = [0]
        for n in layout.numels:
            groups.append(groups[-1] + (n + 7) // 8)
        hit = (torch.tensor(layout.offsets, dtype=torch.int64, device=device),
               torch.tensor(groups, dtype=torch.int64, device=device))
        _group_cache[key] = hit
    return hit


def _payload_of(msg: Any) -> tuple[Any, bool]:
    from ..message import DeltaParameterMessage, ParameterMessage

    if isinstance(msg, DeltaParameterMessage):
        return msg.delta_parameter, True
    if isinstance(msg, ParameterMessage):
        return msg.parameter, False
    return None, False


def _is_quantized(payload: Any) -> bool:
    return isinstance(payload, dict) and "__quantized__" in payload


def _is_arena(payload: Any) -> bool:
    return type(payload).__name__ == "ArenaPayload"


def can_fuse(worker_data: Any, old_parameter_dict: Any) -> bool:
    if not runtime.available():
        return False
    payload, is_delta = _payload_of(worker_data)
    if payload is None or getattr(worker_data, "dataset_size", 0) <= 0:
        return False
    if _is_arena(payload):
        old = as_flat(old_parameter_dict) if old_parameter_dict is not None else None
        return not is_delta or (old is not None and old.flat.is_cuda and old.layout == payload.layout)
    old = as_flat(old_parameter_dict) if old_parameter_dict is not None else None
    if old_parameter_dict is not None and (old is None or not old.flat.is_cuda):
        return False
    if is_delta and old is None:
        return False
    if _is_quantized(payload):
        q = payload["__quantized__"]
        return q.payload.is_cuda and (old is None or q.layout == old.layout)
    flat = as_flat(payload)
    if flat is None:
        # a partial dict (FedOBD stage 1) of views into one flat row is still fusable
        part = getattr(payload, "partial_flat", None)
        return part is not None and old is not None
    return flat.flat.is_cuda and flat.flat.dtype == torch.float32 and (old is None or flat.layout == old.layout)


def materialize(worker_data: Any) -> Any:
    """Undo lazy packing so the generic torch path can consume the message."""
    payload, is_delta = _payload_of(worker_data)
    if _is_arena(payload):
        from ..parallel import context as dist_context

        new = payload.to_local(dist_context.get_context())
        if is_delta:
            worker_data.delta_parameter = new
        else:
            worker_data.parameter = new
        return worker_data
    if _is_quantized(payload):
        q = payload["__quantized__"]
        from ..algorithm.quantization.deterministic import nnadq_dequantize_flat
        from ..algorithm.quantization.stochastic import sq_dequantize_flat

        flat = sq_dequantize_flat(q) if q.kind == "sq" else nnadq_dequantize_flat(q)
        new = FlatParameter(flat, q.layout)
        if is_delta:
            worker_data.delta_parameter = new
        else:
            worker_data.parameter = new
    return worker_data


class PartialFlat(dict):
    """Subset of a flat row: tensor views for the kept names + the row + a presence mask
    (FedOBD stage-1 uploads).  Unlisted tensors are "absent"."""

    def __init__(self, flat: torch.Tensor, layout: ParameterLayout, names: list[str]) -> None:
        views = layout.views(flat)
        super().__init__({n: views[n] for n in names})
        self.partial_flat = flat
        self.layout = layout
        present = torch.zeros(len(layout.names), dtype=torch.uint8)
        for n in names:
            present[layout.index[n]] = 1
        self.present = present

    def __reduce__(self):
        return (dict, (dict(self),))


def _nan_check(flag: torch.Tensor) -> None:
    assert int(flag.item()) == 0, "NaN in aggregated parameters"


def reduce_rows_f32(rows: list[torch.Tensor], weights: list[float], old: torch.Tensor | None, old_scale: float,
                    inv_total: float, out: torch.Tensor | None = None, check_nan: bool = True) -> torch.Tensor:
    """out = old_scale*old + inv_total * sum_k weights[k]*rows[k]  (fp64 accumulation, one launch)."""
    device = rows[0].device
    n = rows[0].numel()
    rows = [r if r.device == device else r.to(device, non_blocking=True) for r in rows]
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=device)
    assert out.is_contiguous() and out.data_ptr() % 16 == 0 and all(r.data_ptr() % 16 == 0 for r in rows), \
        "fedavg_reduce_f32 uses 128-bit accesses: rows and output must be 16-byte aligned"
    row_table = runtime.pointer_table(rows, device)
    dst_table = (ctypes.c_void_p * 1)(out.data_ptr())  # host array: copied into the launch arguments
    w = torch.tensor(weights, dtype=torch.float32).pin_memory().to(device, non_blocking=True)
    flag = torch.zeros(1, dtype=torch.int32, device=device)
    runtime.call("dls_fedavg_reduce_f32", row_table.data_ptr(), w.data_ptr(), len(rows), runtime.ptr(old),
                 float(old_scale), float(inv_total), dst_table, 1, 0, n, None, 1, 0, 0, flag.data_ptr(),
                 runtime.stream_ptr(device))
    STATS["f32"] += 1
    if check_nan:
        _nan_check(flag)
    return out


def fused_aggregate(pending: dict, old_parameter_dict: Any, variant: str) -> FlatParameter:
    msgs = [pending[k] for k in sorted(pending)]
    old = as_flat(old_parameter_dict) if old_parameter_dict is not None else None
    payloads = [_payload_of(m) for m in msgs]
    weights = [float(m.dataset_size) for m in msgs]
    total = sum(weights)
    if all(_is_arena(p) for p, _ in payloads) and variant == "fedavg" and len({d for _, d in payloads}) == 1:
        from ..parallel import context as dist_context

        is_delta = payloads[0][1]
        STATS["p2p"] += 1
        return dist_context.get_context().p2p_reduce(
            [p for p, _ in payloads], weights, old.flat if (is_delta and old is not None) else None,
            1.0 if is_delta else 0.0, 1.0 / total)
    if any(_is_arena(p) for p, _ in payloads):
        for m in msgs:
            materialize(m)
        payloads = [_payload_of(m) for m in msgs]
    plain = all(as_flat(p) is not None for p, _ in payloads) and variant == "fedavg"
    if plain and len({d for _, d in payloads}) == 1:
        is_delta = payloads[0][1]
        rows = [as_flat(p).flat for p, _ in payloads]
        layout = as_flat(payloads[0][0]).layout
        device = old.flat.device if old is not None else rows[0].device
        rows = [r.to(device, non_blocking=True) for r in rows]
        out = reduce_rows_f32(rows, weights, old.flat if (is_delta and old is not None) else None,
                              1.0 if is_delta else 0.0, 1.0 / total)
        return FlatParameter(out, layout)
    return _packed_aggregate(msgs, payloads, weights, old, variant)


def _packed_aggregate(msgs: list, payloads: list, weights: list[float], old: FlatParameter | None,
                      variant: str) -> FlatParameter:
    first = payloads[0][0]
    if _is_quantized(first):
        layout = first["__quantized__"].layout
        device = first["__quantized__"].payload.device
    elif as_flat(first) is not None:
        layout, device = as_flat(first).layout, as_flat(first).flat.device
    else:
        layout, device = first.layout, first.partial_flat.device
    if old is not None:
        device = old.flat.device
    t_off, g_off = layout_tables(layout, device)
    rows = (runtime.DlsPackedRow * len(msgs))()
    keep: list[Any] = []  # keep device tensors alive until the launch is enqueued
    for i, ((payload, is_delta), w) in enumerate(zip(payloads, weights)):
        r = rows[i]
        r.weight, r.is_delta = w, int(is_delta)
        if _is_quantized(payload):
            q = payload["__quantized__"]
            r.kind = 1 if q.kind == "sq" else 2
            r.payload, r.aux = q.payload.data_ptr(), (q.aux.data_ptr() if q.aux.numel() else None)
            r.scale, r.offset = q.scale.data_ptr(), q.offset.data_ptr()
            r.bits, r.byte_offsets = q.bits.data_ptr(), q.byte_offsets.data_ptr()
            keep.append(q)
        elif as_flat(payload) is not None:
            r.kind, r.payload = 0, as_flat(payload).flat.data_ptr()
            keep.append(payload)
        else:
            present = payload.present.to(device, non_blocking=True)
            r.kind, r.payload, r.present = 0, payload.partial_flat.data_ptr(), present.data_ptr()
            keep.extend([payload, present])
    host = torch.frombuffer(bytearray(bytes(rows)), dtype=torch.uint8).pin_memory()
    dev_rows = host.to(device, non_blocking=True)
    out = torch.empty(layout.total, dtype=torch.float32, device=device)
    flag = torch.zeros(1, dtype=torch.int32, device=device)
    runtime.call("dls_fedavg_reduce_packed", dev_rows.data_ptr(), len(msgs), runtime.ptr(old.flat) if old else None,
                 t_off.data_ptr(), g_off.data_ptr(), len(layout.names), int(variant == "dropout"), out.data_ptr(),
                 flag.data_ptr(), runtime.stream_ptr(device))
    STATS["packed"] += 1
    _nan_check(flag)
    del keep
    return FlatParameter(out, layout)


def try_weighted_avg(data_dict: dict, weight_dict: dict) -> Any:
    """Fused path of ``AggregationAlgorithm.weighted_avg`` (Shapley subset models)."""
    if not runtime.available():
        return None
    flats = [as_flat(v.parameter) for v in data_dict.values()]
    if any(f is None or not f.flat.is_cuda or f.flat.dtype != torch.float32 for f in flats):
        return None
    out = reduce_rows_f32([f.flat for f in flats], [float(weight_dict[k]) for k in data_dict], None, 0.0, 1.0)
    return FlatParameter(out, flats[0].layout)


# ---- plain PyTorch statement of the same maths (test oracle) ---------------------------
def oracle_reduce(rows: list[torch.Tensor], weights: list[float], old: torch.Tensor | None, old_scale: float,
                  inv_total: float) -> torch.Tensor:
    acc = torch.zeros_like(rows[0], dtype=torch.float64)
    for r, w in zip(rows, weights):
        acc += r.double() * float(torch.tensor(w, dtype=torch.float32))
    res = acc * inv_total
    if old is not None:
        res = res + float(old_scale) * old.double()
    return res.float()
