"""Star topology across ranks: every rank hosts its share of the workers (threads) plus a
replica of the server.  Local hand-offs work exactly like ``ThreadCentralTopology``; when a
server replica asks for worker data, all replicas first wait for their local workers and
then swap the messages' metadata (gloo ``all_gather_object``).  Flat fp32 payloads never
travel: they are copied once into the rank's peer-mapped upload arena and the metadata
carries ``(rank, slot)``; other tensors are pickled through the control plane, which is
the slow-but-general path (and the one CPU tests exercise with the gloo backend).

Because every replica sees the same messages in the same order and draws from identically
seeded RNGs, the replicas take identical decisions (client selection, early stop, phase
switches) without a leader.
"""
import copy
import dataclasses
from typing import Any

import torch

from ..engine.model_util import as_flat
from ..engine.tensor import recursive_tensor_op
from ..message import DeltaParameterMessage, ParameterMessage
from ..topology.topology import ThreadCentralTopology
from .context import ArenaPayload, DistributedContext


class _ArenaRef:
    def __init__(self, rank: int, slot: int) -> None:
        self.rank, self.slot = rank, slot


class _SignRef:
    """Packed sign words of a signSGD step parked in the owner's peer-mapped buffer."""

    def __init__(self, rank: int, slot: int, parity: int) -> None:
        self.rank, self.slot, self.parity = rank, slot, parity


def _to_cpu(data: Any) -> Any:
    return recursive_tensor_op(data, lambda t: t.detach().cpu())


class DistributedTopology(ThreadCentralTopology):
    def __init__(self, worker_num: int, ctx: DistributedContext) -> None:
        super().__init__(worker_num)
        self.ctx = ctx
        self.local_workers = [w for w in range(worker_num) if ctx.is_local(w)]
        self._inbox: dict[int, Any] = {}
        self.exchanges = 0
        # False (default): every worker reports once per exchange (aggregation servers) -- a replica
        # blocks until all of its local workers have data, then swaps.  True: workers may report at
        # different times (per-step protocols such as sign_SGD) -- replicas swap whatever is there,
        # polling every few milliseconds while nothing is.
        self.partial_exchange = False

    def is_local(self, worker_id: int) -> bool:
        return self.ctx.is_local(worker_id)

    # ---- encoding ---------------------------------------------------------------------------
    def _encode(self, worker_id: int, msg: Any) -> tuple[Any, Any]:
        """Returns (wire object for the peers, message for the local replica)."""
        if msg is None or not dataclasses.is_dataclass(msg):
            return _to_cpu(copy.copy(msg)) if msg is not None else None, msg
        field = "delta_parameter" if isinstance(msg, DeltaParameterMessage) else (
            "parameter" if isinstance(msg, ParameterMessage) else None)
        ctx = self.ctx
        words = msg.other_data.get("gradient") if isinstance(getattr(msg, "other_data", None), dict) else None
        if field is None and ctx.signs is not None and isinstance(words, torch.Tensor) and words.is_cuda \
                and words.dtype == torch.int32 and words.numel() == ctx.sign_words and "parity" in msg.other_data:
            # signSGD step: the words stay on the device; peers read them over NVLink inside their update kernel
            ref = _SignRef(*ctx.store_sign_words(worker_id, int(msg.other_data["parity"]), words))
            wire = copy.copy(msg)
            wire.other_data = {k: (ref if k == "gradient" else _to_cpu(v)) for k, v in msg.other_data.items()}
            return wire, msg
        if field is not None and ctx.arena is not None:
            payload = getattr(msg, field)
            flat = None
            if isinstance(payload, dict) and "__quantized__" in payload and payload["__quantized__"].payload.is_cuda:
                from ..ops.reduce import materialize

                materialize(msg)
                payload = getattr(msg, field)
            fp = as_flat(payload)
            if fp is not None and fp.flat.is_cuda and fp.flat.dtype == torch.float32 and fp.layout == ctx.layout:
                flat = fp.flat
            if flat is not None:
                arena_payload = ctx.store_upload(worker_id, flat)
                local_msg = copy.copy(msg)
                setattr(local_msg, field, arena_payload)
                wire = copy.copy(msg)
                setattr(wire, field, _ArenaRef(ctx.rank, arena_payload.slot))
                wire.other_data = _to_cpu(dict(msg.other_data))
                return wire, local_msg
        wire = copy.copy(msg)
        for f in dataclasses.fields(msg):
            setattr(wire, f.name, _to_cpu(copy.copy(getattr(msg, f.name))))
        return wire, msg

    def _decode(self, wire: Any) -> Any:
        if wire is None or not dataclasses.is_dataclass(wire):
            return wire
        ref = wire.other_data.get("gradient") if isinstance(getattr(wire, "other_data", None), dict) else None
        if isinstance(ref, _SignRef):
            from ..ops.misc import PeerWords

            wire.other_data["gradient"] = PeerWords(self.ctx.sign_words_ptr(ref.rank, ref.slot, ref.parity),
                                                    self.ctx.sign_words, self.ctx.device)
        for name in ("delta_parameter", "parameter"):
            ref = getattr(wire, name, None)
            if isinstance(ref, _ArenaRef):
                setattr(wire, name, self.ctx.remote_payload(ref.rank, ref.slot))
        if self.ctx.device.type == "cuda":
            # generic tensors arrive on the CPU; the algorithms expect them next to the model
            def up(t: torch.Tensor) -> torch.Tensor:
                return t.to(self.ctx.device, non_blocking=True)

            for f in dataclasses.fields(wire):
                v = getattr(wire, f.name)
                if not isinstance(v, ArenaPayload):
                    setattr(wire, f.name, recursive_tensor_op(v, up))
        return wire

    # ---- server side ------------------------------------------------------------------------------
    def server_has_data(self, worker_id: int) -> bool:
        self._check()
        return worker_id in self._inbox

    def server_get(self, worker_id: int) -> Any:
        return self._inbox.pop(worker_id)

    def server_wait(self, worker_ids: set[int], timeout: float | None) -> set[int]:
        have = {w for w in worker_ids if w in self._inbox}
        if have:
            return have
        local_needed = sorted(w for w in worker_ids if self.is_local(w))
        if self.partial_exchange:
            return self._partial_wait(worker_ids, local_needed, timeout)
        local_msgs: dict[int, Any] = {}
        for w in local_needed:
            ready = super().server_wait({w}, timeout)
            if not ready:
                return set()
            local_msgs[w] = super().server_get(w)
        wire: dict[int, Any] = {}
        for w, msg in local_msgs.items():
            wire[w], local_msgs[w] = self._encode(w, msg)
        self.ctx.flush_sign_stores()
        gathered = self.ctx.exchange(wire)
        self.exchanges += 1
        for rank, part in enumerate(gathered):
            for w, obj in part.items():
                self._inbox[w] = local_msgs[w] if rank == self.ctx.rank else self._decode(obj)
        return {w for w in worker_ids if w in self._inbox}

    def _partial_wait(self, worker_ids: set[int], local_needed: list[int], timeout: float | None) -> set[int]:
        import time

        deadline = None if timeout is None else time.monotonic() + timeout
        while True:
            ready = super().server_wait(set(local_needed), 0.003) if local_needed else set()
            if not local_needed:
                time.sleep(0.003)
            local_msgs = {w: super().server_get(w) for w in sorted(ready)}
            wire: dict[int, Any] = {}
            for w, msg in local_msgs.items():
                wire[w], local_msgs[w] = self._encode(w, msg)
            self.ctx.flush_sign_stores()  # the words must be in memory before a peer learns about them
            gathered = self.ctx.exchange(wire)
            self.exchanges += 1
            for rank, part in enumerate(gathered):
                for w, obj in part.items():
                    self._inbox[w] = local_msgs[w] if rank == self.ctx.rank else self._decode(obj)
            have = {w for w in worker_ids if w in self._inbox}
            # every replica sees the same inbox, so they all leave the loop in the same iteration
            if have:
                return have
            if deadline is not None and time.monotonic() > deadline:
                return set()

    def server_send(self, worker_id: int, data: Any) -> None:
        if self.is_local(worker_id):
            super().server_send(worker_id, data)

    def pre_send_fence(self) -> None:
        if self.ctx.peer_reads_pending:
            self.ctx.fence()
