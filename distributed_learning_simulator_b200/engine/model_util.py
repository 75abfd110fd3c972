"""Model introspection + the flat parameter arena.

Own version of ``cyy_torch_toolbox.model.util.ModelUtil`` with the contract the
reference uses (/root/reference/simulation_lib/method/fed_obd/obd_algorithm.py:9-22,33-86,
server/server.py:47-48, util/model.py:15,23, worker/graph_worker.py:331).

B200-first addition: ``flatten_parameters()`` re-homes every parameter (and its
gradient) as a view into ONE contiguous fp32 buffer.  Every communication hot path
(FedAvg reduction, quantizers, sign packing, block scores, the fused SGD step) is then a
1-D kernel over ``[P]`` with a per-tensor segment table, and a worker upload is a
``[P]`` row of a device arena instead of a pickled dict of CPU tensors.
"""
import os
from typing import Any, Iterable

import torch
from torch import nn

from .tensor import TensorDict


def _is_channels_last(t: torch.Tensor) -> bool:
    """4-D tensor physically stored as N,H,W,C (and not also plain-contiguous)."""
    return t.dim() == 4 and not t.is_contiguous() and t.is_contiguous(memory_format=torch.channels_last)


class ParameterLayout:
    """Segment table of a flat parameter vector: names, shapes, offsets, and which 4-D
    tensors are stored channels-last (N,H,W,C order inside their segment) so that cuDNN's
    tensor-core convolutions need no layout conversion kernels."""

    def __init__(self, names: list[str], shapes: list[torch.Size], formats: list[int] | None = None) -> None:
        self.names = list(names)
        self.shapes = [torch.Size(s) for s in shapes]
        self.formats = list(formats) if formats is not None else [0] * len(self.names)
        self.numels = [int(s.numel()) for s in self.shapes]
        self.offsets = [0]
        for n in self.numels:
            self.offsets.append(self.offsets[-1] + n)
        self.total = self.offsets[-1]
        self.index = {n: i for i, n in enumerate(self.names)}

    @classmethod
    def from_dict(cls, parameter_dict: TensorDict) -> "ParameterLayout":
        return cls(list(parameter_dict.keys()), [v.shape for v in parameter_dict.values()],
                   [int(_is_channels_last(v)) for v in parameter_dict.values()])

    def __eq__(self, other: object) -> bool:
        return isinstance(other, ParameterLayout) and self.names == other.names and self.shapes == other.shapes \
            and self.formats == other.formats

    def __hash__(self) -> int:
        return hash((tuple(self.names), tuple(tuple(s) for s in self.shapes), tuple(self.formats)))

    def slice_of(self, name: str) -> slice:
        i = self.index[name]
        return slice(self.offsets[i], self.offsets[i + 1])

    def view_of(self, flat: torch.Tensor, i: int) -> torch.Tensor:
        seg = flat[self.offsets[i] : self.offsets[i + 1]]
        shape = self.shapes[i]
        if self.formats[i]:
            o, c, h, w = shape
            return seg.view(o, h, w, c).permute(0, 3, 1, 2)
        return seg.view(shape)

    def views(self, flat: torch.Tensor) -> TensorDict:
        assert flat.numel() == self.total
        return {n: self.view_of(flat, i) for i, n in enumerate(self.names)}

    def flatten(self, parameter_dict: TensorDict, out: torch.Tensor | None = None) -> torch.Tensor:
        if out is None:
            ref = next(iter(parameter_dict.values()))
            out = torch.empty(self.total, dtype=torch.float32, device=ref.device)
        for i, n in enumerate(self.names):
            self.view_of(out, i).copy_(parameter_dict[n])
        return out

    def segment_offsets(self, device: torch.device | str = "cpu") -> torch.Tensor:
        return torch.tensor(self.offsets, dtype=torch.int64, device=device)


class FlatParameter(dict):
    """A TensorDict whose tensors are views of one flat ``[P]`` vector.

    Behaves as a plain ``dict[str, Tensor]``; ``.flat``/``.layout`` let the fused
    kernels skip the per-tensor walk."""

    def __init__(self, flat: torch.Tensor, layout: ParameterLayout) -> None:
        super().__init__(layout.views(flat))
        self.flat = flat
        self.layout = layout

    def is_intact(self) -> bool:
        """False once someone replaced/added/removed entries (then only the dict is valid)."""
        if len(self) != len(self.layout.names):
            return False
        base = self.flat.untyped_storage().data_ptr()
        return all(k in self and self[k].untyped_storage().data_ptr() == base for k in self.layout.names)

    def __reduce__(self):  # pickles as a plain dict of CPU-able tensors
        return (dict, (dict(self),))

    def __copy__(self):  # a shallow copy stays flat (``copy.copy`` would otherwise go through __reduce__)
        if self.is_intact():
            return FlatParameter(self.flat, self.layout)
        return dict(self)


def as_flat(parameter: TensorDict) -> "FlatParameter | None":
    if isinstance(parameter, FlatParameter) and parameter.is_intact():
        return parameter
    return None


class ModelUtil:
    def __init__(self, model: nn.Module) -> None:
        self.__model = model
        self.__flat: torch.Tensor | None = None
        self.__flat_grad: torch.Tensor | None = None
        self.__layout: ParameterLayout | None = None
        self.__shadow: torch.Tensor | None = None  # bf16 compute copy of the arena (see enable_shadow)
        self.__shadow_store: torch.Tensor | None = None  # kept across disable/enable: CUDA graphs hold its address

    @property
    def model(self) -> nn.Module:
        return self.__model

    # ---- parameters -----------------------------------------------------------
    def get_parameter_dict(self, detach: bool = True) -> TensorDict:
        if self.__shadow is not None:
            # the model's matrices are bf16 compute copies: the parameters ARE the fp32 master arena
            return self.layout.views(self.__flat)
        res = {}
        for name, p in self.__model.named_parameters():
            res[name] = p.detach() if detach else p
        return res

    def get_parameter_list(self, detach: bool = True) -> torch.Tensor:
        if self.__flat is not None:
            return self.__flat.detach() if detach else self.__flat
        return torch.cat([p.detach().reshape(-1) for p in self.__model.parameters()])

    @property
    def layout(self) -> ParameterLayout:
        if self.__layout is None:
            self.__layout = ParameterLayout.from_dict(self.get_parameter_dict())
        return self.__layout

    def load_parameter_dict(self, parameter_dict: TensorDict, check_parameter: bool = False) -> None:
        own = self.get_parameter_dict() if self.__shadow is not None else dict(self.__model.named_parameters())
        if check_parameter:
            assert set(own) == set(parameter_dict), "parameter names differ"
        with torch.no_grad():
            for name, value in parameter_dict.items():
                own[name].copy_(value.to(own[name].device, non_blocking=True))
        self.refresh_shadow()

    def load_parameter_list(self, flat: torch.Tensor) -> None:
        with torch.no_grad():
            if self.__flat is not None:
                self.__flat.copy_(flat.to(self.__flat.device, non_blocking=True))
                self.refresh_shadow()
                return
            off = 0
            for p in self.__model.parameters():
                n = p.numel()
                p.copy_(flat[off : off + n].view_as(p))
                off += n

    # ---- flat arena -------------------------------------------------------------
    def flatten_parameters(self, with_grad: bool = False, channels_last: bool | None = None) -> torch.Tensor:
        """Make every parameter a view of one flat fp32 buffer (idempotent per device).
        On CUDA, 4-D (convolution) weights are stored channels-last inside their segment."""
        params = list(self.__model.parameters())
        assert params
        device = params[0].device
        if self.__flat is not None and self.__flat.device == device and self.is_flat():
            return self.__flat
        if channels_last is None:
            channels_last = device.type == "cuda" and os.environ.get("DLS_CHANNELS_LAST", "1") != "0"
        total = sum(p.numel() for p in params)
        flat = torch.empty(total, dtype=torch.float32, device=device)
        flat_grad = torch.zeros(total, dtype=torch.float32, device=device) if with_grad else None
        off = 0
        with torch.no_grad():
            for p in params:
                n = p.numel()
                seg = flat[off : off + n]
                if channels_last and p.dim() == 4 and p.shape[1] > 1 and (p.shape[2] > 1 or p.shape[3] > 1):
                    o, c, h, w = p.shape
                    view = seg.view(o, h, w, c).permute(0, 3, 1, 2)
                else:
                    view = seg.view(p.shape)
                view.copy_(p.data)
                p.data = view
                if flat_grad is not None and p.requires_grad:
                    p.grad = flat_grad[off : off + n].view(p.shape)
                off += n
        self.__flat, self.__flat_grad = flat, flat_grad
        self.__layout = None
        self.__shadow_store = None
        return flat

    def is_flat(self) -> bool:
        if self.__flat is None:
            return False
        base = self.__flat.untyped_storage().data_ptr()
        if self.__shadow is not None:
            sbase = self.__shadow.untyped_storage().data_ptr()
            return all(p.data.untyped_storage().data_ptr() in (base, sbase) for p in self.__model.parameters())
        return all(p.data.untyped_storage().data_ptr() == base for p in self.__model.parameters())

    # ---- bf16 compute shadow ---------------------------------------------------------
    @property
    def shadow(self) -> torch.Tensor | None:
        return self.__shadow

    def shadow_flags(self) -> list[int]:
        """1 for tensors whose model parameter is the bf16 shadow copy: matrices / filters whose
        shadow segment starts on a 16-byte boundary (cuDNN/cuBLAS pick their vectorised kernels by
        pointer alignment; an arena offset that is fine for fp32 can be misaligned at 2 bytes/element)."""
        layout = self.layout
        return [int(len(shape) >= 2 and off % 8 == 0) for shape, off in zip(layout.shapes, layout.offsets)]

    def enable_shadow(self) -> torch.Tensor:
        """Mixed-precision training without per-step cast kernels: every >= 2-D parameter of the
        model becomes a bf16 view into a second ("shadow") arena with the same segment layout, while
        the fp32 arena stays the master copy that is exchanged, check-pointed and updated by the
        fused SGD kernel (which rewrites the shadow in the same pass).  1-D parameters (BatchNorm
        scale/shift, biases) stay fp32 views of the master."""
        assert self.__flat is not None and self.is_flat()
        if self.__shadow is not None:
            return self.__shadow
        layout = self.layout
        shadow = self.__shadow_store
        if shadow is None or shadow.device != self.__flat.device or shadow.numel() != layout.total:
            shadow = torch.empty(layout.total, dtype=torch.bfloat16, device=self.__flat.device)
        self.__shadow_store = shadow
        shadow.copy_(self.__flat)
        params = dict(self.__model.named_parameters())
        flags = self.shadow_flags()
        for i, name in enumerate(layout.names):
            if flags[i]:
                p = params[name]
                p.data = layout.view_of(shadow, i)
                p.grad = None
        self.__shadow = shadow
        return shadow

    def disable_shadow(self) -> None:
        if self.__shadow is None:
            return
        layout = self.layout
        params = dict(self.__model.named_parameters())
        self.__shadow = None
        for i, name in enumerate(layout.names):
            p = params[name]
            p.data = layout.view_of(self.__flat, i)
            p.grad = None

    def refresh_shadow(self) -> None:
        if self.__shadow is not None:
            with torch.no_grad():
                self.__shadow.copy_(self.__flat)

    @property
    def flat_parameter(self) -> torch.Tensor | None:
        return self.__flat if self.is_flat() else None

    @property
    def flat_gradient(self) -> torch.Tensor | None:
        return self.__flat_grad if self.is_flat() else None

    def rebind_flat_gradients(self) -> None:
        """Point ``p.grad`` back at the flat gradient buffer (after zero_grad(set_to_none))."""
        if self.__flat_grad is None:
            return
        off = 0
        for p in self.__model.parameters():
            n = p.numel()
            if p.requires_grad:
                p.grad = self.__flat_grad[off : off + n].view(p.shape)
            off += n

    def to_device(self, device: torch.device) -> None:
        first = next(self.__model.parameters(), None)
        if first is not None and first.device == device:
            return
        self.disable_shadow()
        self.__model.to(device)
        self.__flat = None
        self.__flat_grad = None

    # ---- modules ----------------------------------------------------------------
    def get_modules(self) -> list[tuple[str, nn.Module]]:
        return list(self.__model.named_modules())

    def get_sub_modules(self) -> list[tuple[str, nn.Module]]:
        """Leaf modules in definition order."""
        return [(n, m) for n, m in self.__model.named_modules() if not list(m.children())]

    def have_module(self, module_type: type | None = None, module_name: str | None = None) -> bool:
        for name, m in self.__model.named_modules():
            if module_type is not None and isinstance(m, module_type):
                return True
            if module_name is not None and name == module_name:
                return True
        return False

    def get_module_blocks(self, block_types: Iterable[tuple]) -> list[list[tuple[str, nn.Module]]]:
        """Find runs of consecutive sub-modules matching a block type.

        A block type is a tuple of classes or class-name strings.  Single-element
        types match a (possibly non-leaf) module as a whole; its descendants are then
        skipped.  Longer types match consecutive *leaf* modules.  Longest types win.
        """

        def matches(m: nn.Module, t: Any) -> bool:
            if isinstance(t, str):
                return type(m).__name__ == t
            return isinstance(m, t)

        block_types = sorted(set(block_types), key=len, reverse=True)
        blocks: list[list[tuple[str, nn.Module]]] = []
        covered_prefixes: list[str] = []

        def covered(name: str) -> bool:
            return any(name == c or name.startswith(c + ".") for c in covered_prefixes)

        # whole-module (single element) types first: they may be containers
        singles = [t for t in block_types if len(t) == 1]
        for name, m in self.__model.named_modules():
            if not name or covered(name):
                continue
            if any(matches(m, t[0]) for t in singles) and list(m.parameters()):
                if list(m.children()):
                    blocks.append([(name, m)])
                    covered_prefixes.append(name)
        leaves = [(n, m) for n, m in self.get_sub_modules() if not covered(n)]
        i = 0
        multi = [t for t in block_types if len(t) > 1]
        while i < len(leaves):
            hit = None
            for t in multi:
                if i + len(t) <= len(leaves) and all(matches(leaves[i + j][1], t[j]) for j in range(len(t))):
                    hit = t
                    break
            if hit is not None:
                block = leaves[i : i + len(hit)]
                if any(list(m.parameters()) for _, m in block):
                    blocks.append(block)
                i += len(hit)
                continue
            n, m = leaves[i]
            if any(matches(m, t[0]) for t in singles) and list(m.parameters()):
                blocks.append([(n, m)])
            i += 1
        return blocks

    # ---- normalisation layers -----------------------------------------------------
    def disable_running_stats(self) -> None:
        """BatchNorm always normalises with batch statistics; running buffers are
        dropped so they are neither exchanged nor part of the model state
        (reference: util/model.py:23, server/server.py:48)."""
        for m in self.__model.modules():
            if isinstance(m, nn.modules.batchnorm._BatchNorm):
                m.track_running_stats = False
                m.register_buffer("running_mean", None)
                m.register_buffer("running_var", None)
                m.register_buffer("num_batches_tracked", None)
