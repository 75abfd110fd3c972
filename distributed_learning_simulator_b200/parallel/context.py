"""Per-process distributed context: one rank per GPU under ``torch.distributed``.

NCCL (via torch.distributed) is the plumbing: rendezvous, tiny metric all-reduces, the
gloo side group for pickled control messages.  The parameter hot path does not use it:
uploads live in peer-mapped arenas (``csrc/ipc.cu``) and ``p2p_reduce`` launches the
hand-written reduce kernel, in which every rank reduces its own ``[lo, hi)`` slice by
loading all ranks' rows over NVLink and stores the result into every rank's global-model
buffer, with device-side flag barriers on both sides (``csrc/reduce.cu``).
"""
import ctypes
import os
import threading
from typing import Any

import torch
import torch.distributed as dist

from ..engine.log import get_logger
from ..engine.model_util import FlatParameter, ParameterLayout
from ..ops import runtime

_lock = threading.Lock()
_current: "DistributedContext | None" = None


def is_distributed() -> bool:
    return int(os.environ.get("WORLD_SIZE", "1")) > 1 and os.environ.get("DLS_DISABLE_DISTRIBUTED", "0") != "1"


def current() -> "DistributedContext | None":
    return _current


def get_context() -> "DistributedContext":
    global _current
    with _lock:
        if _current is None:
            _current = DistributedContext()
    return _current


class SymmetricBuffer:
    """``nbytes`` of zeroed device memory on every rank, each mapped into every process."""

    def __init__(self, ctx: "DistributedContext", nbytes: int) -> None:
        self.nbytes = int(nbytes)
        self.ctx = ctx
        lib = runtime.require()
        local = ctypes.c_void_p()
        rc = lib.dls_ipc_alloc(self.nbytes, ctypes.byref(local))
        if rc != 0:
            raise RuntimeError(lib.dls_last_error().decode())
        self.local_ptr = int(local.value)
        handle = ctypes.create_string_buffer(64)
        rc = lib.dls_ipc_get_handle(self.local_ptr, handle)
        if rc != 0:
            raise RuntimeError(lib.dls_last_error().decode())
        handles: list[Any] = [None] * ctx.world
        dist.all_gather_object(handles, bytes(handle.raw), group=ctx.control)
        self.ptrs: list[int] = []
        for r, h in enumerate(handles):
            if r == ctx.rank:
                self.ptrs.append(self.local_ptr)
                continue
            peer = ctypes.c_void_p()
            rc = lib.dls_ipc_open_handle(h, ctypes.byref(peer))
            if rc != 0:
                raise RuntimeError(f"cannot map rank {r}'s buffer: {lib.dls_last_error().decode()}")
            self.ptrs.append(int(peer.value))
        dist.barrier(group=ctx.control)

    def local_tensor(self, dtype: torch.dtype, numel: int, byte_offset: int = 0) -> torch.Tensor:
        itemsize = torch.empty(0, dtype=dtype).element_size()
        assert byte_offset + numel * itemsize <= self.nbytes

        class _Raw:
            pass

        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (numel * itemsize,), "typestr": "|u1",
                                        "data": (self.local_ptr + byte_offset, False), "version": 2}
        raw._keepalive = self
        t = torch.as_tensor(raw, device=self.ctx.device)
        return t.view(dtype)

    def ptr(self, rank: int, byte_offset: int = 0) -> int:
        return self.ptrs[rank] + byte_offset


class MulticastBuffer:
    """Same contract as ``SymmetricBuffer`` plus ``mc_ptr``: one virtual address whose stores the
    NVSwitch replicates into every rank's copy (NVLS).  Allocation, handle exchange and the multicast
    binding are PyTorch's symmetric-memory plumbing; the kernels that write through ``mc_ptr`` are
    ours (csrc/reduce.cu: ``multimem.st``).  Raises when the platform has no multicast support."""

    def __init__(self, ctx: "DistributedContext", nbytes: int) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        self.nbytes = int(nbytes)
        self.ctx = ctx
        self._tensor = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=ctx.device)
        self._tensor.zero_()
        self._handle = symm_mem.rendezvous(self._tensor, dist.group.WORLD)
        self.mc_ptr = int(getattr(self._handle, "multicast_ptr", 0) or 0)
        if not self.mc_ptr:
            raise RuntimeError("symmetric memory has no multicast address on this platform")
        self.ptrs = [int(p) for p in self._handle.buffer_ptrs]
        self.local_ptr = self.ptrs[ctx.rank]
        assert self.local_ptr == self._tensor.data_ptr()
        torch.cuda.synchronize(ctx.device)
        dist.barrier(group=ctx.control)

    def local_tensor(self, dtype: torch.dtype, numel: int, byte_offset: int = 0) -> torch.Tensor:
        itemsize = torch.empty(0, dtype=dtype).element_size()
        assert byte_offset + numel * itemsize <= self.nbytes and byte_offset % itemsize == 0
        return self._tensor[byte_offset: byte_offset + numel * itemsize].view(dtype)

    def ptr(self, rank: int, byte_offset: int = 0) -> int:
        return self.ptrs[rank] + byte_offset


class ArenaPayload(dict):
    """Message payload that lives in a rank's upload arena.  ``ptr`` is valid in THIS
    process (local or peer-mapped); ``local`` is a tensor view when the row is ours."""

    def __init__(self, rank: int, slot: int, ptr: int, layout: ParameterLayout, local: torch.Tensor | None) -> None:
        super().__init__()
        self.rank, self.slot, self.ptr, self.layout, self.local = rank, slot, ptr, layout, local

    def to_local(self, ctx: "DistributedContext") -> FlatParameter:
        if self.local is not None:
            return FlatParameter(self.local.clone(), self.layout)
        out = torch.empty(self.layout.total, dtype=torch.float32, device=ctx.device)
        runtime.call("dls_copy_device", out.data_ptr(), self.ptr, self.layout.total * 4, runtime.stream_ptr(ctx.device))
        ctx.peer_reads_pending = True
        return FlatParameter(out, self.layout)


class DistributedContext:
    def __init__(self) -> None:
        self._verdict_host: torch.Tensor | None = None
        self._verdict_event: Any = None
        self._verdict_pending = False
        self.two_stage_reductions = 0
        self._agg_checks_left = int(os.environ.get("DLS_AGG_CHECK", "0"))
        self.agg_max_abs_err = 0.0
        self.agg_checks_done = 0
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        use_cuda = torch.cuda.is_available() and os.environ.get("DLS_FORCE_CPU", "0") != "1"
        self.device = torch.device("cuda", self.local_rank) if use_cuda else torch.device("cpu")
        if use_cuda:
            torch.cuda.set_device(self.device)
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if use_cuda:
                dist.init_process_group("nccl", device_id=self.device)
            else:
                dist.init_process_group("gloo")
        self.control = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else dist.group.WORLD
        self.use_p2p = use_cuda and runtime.available() and os.environ.get("DLS_DISABLE_P2P", "0") != "1"
        self.layout: ParameterLayout | None = None
        self.arena: SymmetricBuffer | None = None
        self.globals: SymmetricBuffer | None = None
        self.flags: SymmetricBuffer | None = None
        self.signs: SymmetricBuffer | None = None
        self.sign_words = 0
        self.sign_stores_pending = False
        self.slots = 0
        self.epoch = 0
        self.global_index = 0
        self.peer_reads_pending = False
        self._flag_table: torch.Tensor | None = None
        self.p2p_reductions = 0
        self.multicast = False
        self.multicast_reduce = False
        self.multicast_reductions = 0
        self._table_cache: dict = {}
        get_logger().debug("rank %s/%s on %s, p2p=%s", self.rank, self.world, self.device, self.use_p2p)

    # ---- ownership --------------------------------------------------------------------
    def owner_of(self, worker_id: int) -> int:
        return worker_id % self.world

    def is_local(self, worker_id: int) -> bool:
        return self.owner_of(worker_id) == self.rank

    def slot_of(self, worker_id: int) -> int:
        return worker_id // self.world

    # ---- symmetric memory -----------------------------------------------------------------
    def setup_arena(self, layout: ParameterLayout, worker_number: int) -> None:
        """Collective.  One upload row per local worker, two global-model buffers, flag words."""
        if not self.use_p2p:
            return
        if self.layout is not None and self.layout == layout and self.slots >= -(-worker_number // self.world):
            return
        self.layout = layout
        self.slots = -(-worker_number // self.world)
        self.row_bytes = ((layout.total * 4 + 255) // 256) * 256
        # slots upload rows + one row for this rank's pre-reduced partial (two-stage reduction)
        self.arena = SymmetricBuffer(self, self.row_bytes * (self.slots + 1))
        self.globals = self._global_buffers(self.row_bytes * 2)
        self.flags = SymmetricBuffer(self, 256)
        # signSGD: two generations (step parity) of packed sign words per local worker, read by the peers' fused
        # vote + update kernels; a worker is at most one step ahead of the slowest reader
        self.sign_words = (layout.total + 31) // 32
        self.sign_row_bytes = ((self.sign_words * 4 + 255) // 256) * 256
        self.signs = SymmetricBuffer(self, 2 * self.slots * self.sign_row_bytes)
        self.sign_stores_pending = False
        self._flag_table = runtime.pointer_table(self.flags.ptrs, self.device)
        self._nan = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.epoch = 0

    def _global_buffers(self, nbytes: int) -> Any:
        """The two global-model buffers: multicast-bound when NVLS is available on every rank
        (``DLS_NVLS=0`` forces the per-peer store path), plain IPC-mapped otherwise."""
        buf: Any = None
        if os.environ.get("DLS_NVLS", "1") != "0":
            try:
                buf = MulticastBuffer(self, nbytes)
            except Exception as e:  # noqa: BLE001 - no NVSwitch / no fabric handles / old driver
                get_logger().info("NVLS multicast unavailable on rank %d: %s", self.rank, e)
                buf = None
        everyone = all(self.exchange(buf is not None))
        if buf is not None and not everyone:
            buf = None
        self.multicast = buf is not None
        # Measured (profiles/agg_sweep.md): in the all-to-all pattern of the fused reduce every GPU's NVLink
        # INGRESS carries the peers' rows plus the peers' slices either way, and a multicast store also sends
        # the writer's own slice back through the switch, so per-destination peer stores are the faster
        # broadcast half there; the multicast store wins for one-to-all broadcasts (dls_multimem_bcast_f32).
        # DLS_NVLS_REDUCE=1 selects the multimem.st epilogue in the fused kernel as well.
        self.multicast_reduce = self.multicast and os.environ.get("DLS_NVLS_REDUCE", "0") == "1"
        return buf if buf is not None else SymmetricBuffer(self, nbytes)

    def arena_row(self, slot: int) -> torch.Tensor:
        assert self.arena is not None and self.layout is not None
        return self.arena.local_tensor(torch.float32, self.layout.total, slot * self.row_bytes)

    def store_upload(self, worker_id: int, flat: torch.Tensor) -> ArenaPayload:
        slot = self.slot_of(worker_id)
        row = self.arena_row(slot)
        row.copy_(flat, non_blocking=True)
        return ArenaPayload(self.rank, slot, self.arena.ptr(self.rank, slot * self.row_bytes), self.layout, row)

    def _sign_offset(self, slot: int, parity: int) -> int:
        return ((parity & 1) * self.slots + slot) * self.sign_row_bytes

    def store_sign_words(self, worker_id: int, parity: int, words: torch.Tensor) -> tuple[int, int, int]:
        """Copy a local worker's packed sign words into the peer-mapped buffer (current stream); the caller
        synchronises before announcing them (``flush_sign_stores``)."""
        assert self.signs is not None and words.numel() == self.sign_words
        slot = self.slot_of(worker_id)
        self.signs.local_tensor(torch.int32, self.sign_words, self._sign_offset(slot, parity)).copy_(words, non_blocking=True)
        self.sign_stores_pending = True
        return self.rank, slot, parity & 1

    def flush_sign_stores(self) -> None:
        if self.sign_stores_pending:
            torch.cuda.current_stream(self.device).synchronize()
            self.sign_stores_pending = False

    def sign_words_ptr(self, rank: int, slot: int, parity: int) -> int:
        assert self.signs is not None
        return self.signs.ptr(rank, self._sign_offset(slot, parity))

    def remote_payload(self, rank: int, slot: int) -> ArenaPayload:
        assert self.arena is not None and self.layout is not None
        local = self.arena_row(slot) if rank == self.rank else None
        return ArenaPayload(rank, slot, self.arena.ptr(rank, slot * self.row_bytes), self.layout, local)

    # ---- the fused reduce-scatter + all-gather ------------------------------------------------
    def p2p_reduce(self, payloads: list[ArenaPayload], weights: list[float], old: torch.Tensor | None,
                   old_scale: float, inv_total: float) -> FlatParameter:
        assert self.globals is not None and self.layout is not None and self.flags is not None
        n = self.layout.total
        self._check_previous_verdict()
        all_payloads, all_weights = payloads, weights
        if len(payloads) > self.world and os.environ.get("DLS_TWO_STAGE", "1") != "0":
            # Two-stage: many workers per GPU.  Stage 1 folds this rank's own rows into ONE weighted fp32 partial
            # (local HBM traffic only); stage 2 is the cross-GPU kernel over the G partials, so NVLink carries
            # (G - 1) / G of a row per GPU instead of a share of every remote worker's row.
            from ..ops.reduce import reduce_rows_f32

            mine = [(p.local, w) for p, w in zip(payloads, weights) if p.rank == self.rank and p.local is not None]
            partial = self.arena_row(self.slots)
            if mine:
                reduce_rows_f32([r for r, _ in mine], [w for _, w in mine], None, 0.0, 1.0, out=partial, check_nan=False)
            else:
                partial.zero_()
            self.two_stage_reductions += 1
            payloads = [ArenaPayload(r, self.slots, self.arena.ptr(r, self.slots * self.row_bytes), self.layout,
                                     partial if r == self.rank else None) for r in range(self.world)]
            weights = [1.0] * self.world
        self.global_index ^= 1
        off = self.global_index * self.row_bytes
        if old is not None and old.data_ptr() == self.globals.local_ptr + off:
            raise RuntimeError("old model aliases the destination buffer")
        chunk = ((-(-n // self.world)) + 3) // 4 * 4
        lo, hi = min(self.rank * chunk, n), min((self.rank + 1) * chunk, n)
        # the same arena slots and dataset sizes come back every round: keep the tiny device tables
        key = (tuple(p.ptr for p in payloads), tuple(weights))
        cached = self._table_cache.get(key)
        if cached is None:
            if len(self._table_cache) > 64:
                self._table_cache.clear()
            cached = (runtime.pointer_table([p.ptr for p in payloads], self.device),
                      torch.tensor(weights, dtype=torch.float32).pin_memory().to(self.device, non_blocking=True))
            self._table_cache[key] = cached
        rows, w = cached
        self.epoch += 1
        self._nan.zero_()
        if self.multicast_reduce:
            # NVLS: each output vector is one multimem.st that the switch replicates into every rank's buffer
            runtime.call("dls_fedavg_reduce_f32_mc", rows.data_ptr(), w.data_ptr(), len(payloads), runtime.ptr(old),
                         float(old_scale), float(inv_total), self.globals.mc_ptr + off, lo, hi,
                         self._flag_table.data_ptr(), self.world, self.rank, self.epoch, self._nan.data_ptr(),
                         runtime.stream_ptr(self.device))
            self.multicast_reductions += 1
        else:
            dsts = (ctypes.c_void_p * self.world)(*[self.globals.ptr(r, off) for r in range(self.world)])
            runtime.call("dls_fedavg_reduce_f32", rows.data_ptr(), w.data_ptr(), len(payloads), runtime.ptr(old),
                         float(old_scale), float(inv_total), dsts, self.world, lo, hi, self._flag_table.data_ptr(),
                         self.world, self.rank, self.epoch, self._nan.data_ptr(), runtime.stream_ptr(self.device))
        self.p2p_reductions += 1
        if self._agg_checks_left > 0:
            # self-check (bench warm-up rounds, DLS_AGG_CHECK=<calls>): the same reduction as an fp64 NCCL
            # all-reduce of every rank's locally weighted rows
            self._agg_checks_left -= 1
            acc = torch.zeros(n, dtype=torch.float64, device=self.device)
            for payload, wt in zip(all_payloads, all_weights):
                if payload.rank == self.rank and payload.local is not None:
                    acc.add_(payload.local.double(), alpha=float(wt))
            dist.all_reduce(acc)
            acc.mul_(float(inv_total))
            if old is not None:
                acc.add_(old.double(), alpha=float(old_scale))
            got = self.globals.local_tensor(torch.float32, n, off).double()
            self.agg_max_abs_err = max(self.agg_max_abs_err, float((got - acc).abs().max().item()))
            self.agg_checks_done += 1
        # the kernel's verdict words (NaN per rank, barrier time-out) are read back without stalling the round:
        # an asynchronous copy now, the check when the next reduction starts (or at finish_reductions())
        if self._verdict_host is None:
            self._verdict_host = torch.zeros(9, dtype=torch.int32).pin_memory()
            self._verdict_event = torch.cuda.Event()
        self._verdict_host.copy_(self.flags.local_tensor(torch.int32, 32)[16:25], non_blocking=True)
        self._verdict_event.record(torch.cuda.current_stream(self.device))
        self._verdict_pending = True
        if os.environ.get("DLS_SYNC_VERDICT", "0") == "1":
            self._check_previous_verdict()
        return FlatParameter(self.globals.local_tensor(torch.float32, n, off), self.layout)

    def _check_previous_verdict(self) -> None:
        if not self._verdict_pending:
            return
        self._verdict_pending = False
        self._verdict_event.synchronize()
        verdict = self._verdict_host.tolist()
        if verdict[8]:
            raise RuntimeError("peer barrier timed out inside fedavg_reduce (a rank is missing)")
        assert not any(verdict[: self.world]), "NaN in aggregated parameters"

    def finish_reductions(self) -> None:
        """Check the verdict of the last fused reduction (called when a run ends)."""
        self._check_previous_verdict()

    # ---- small collectives -------------------------------------------------------------------
    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def exchange(self, obj: Any) -> list[Any]:
        out: list[Any] = [None] * self.world
        dist.all_gather_object(out, obj, group=self.control)
        return out

    def fence(self) -> None:
        """Everyone has finished reading everyone's arena rows (non-fused consumers)."""
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
        dist.barrier(group=self.control)
        self.peer_reads_pending = False
