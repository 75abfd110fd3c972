"""Worker for synchronous per-step gradient exchange (parameter-server SGD).

Role parity with the reference scaffold /root/reference/simulation_lib/worker/
gradient_worker.py:13-131: an ``OPTIMIZER_STEP`` hook replaces the local optimizer step --
gradients (with weight decay folded in, :13-26) are flattened, passed through the
``_process_gradient`` override point (:47-48), sent to the server, the answer is applied
with momentum / dampening / nesterov by hand (:96-116); ``AFTER_EPOCH`` writes
``epoch_stat.json`` (:118-131); ``AFTER_EXECUTE`` reports the end (:44-45).

Differences by design: typed messages instead of the legacy dicts; the
flat gradient is built straight from the arena layout; the update is one fused kernel on
CUDA (``dls_sgd_step`` with a one-entry gradient table) instead of a Python loop over
parameter tensors.
"""
import json
import os
from typing import Any

import torch

from ..engine.ml_type import ExecutorHookPoint
from ..message import Message
from ..ops import runtime
from ..ops import sgd as sgd_ops
from .client import Client


def compute_gradient(params: list, d_p_list: list, weight_decay: float) -> list:
    real_gradient = []
    for param, d_p in zip(params, d_p_list):
        if weight_decay != 0:
            d_p = d_p.add(param, alpha=weight_decay)
        real_gradient.append(d_p)
    return real_gradient


class GradientMessage(Message):
    """Per-step upload/download: ``other_data["gradient"]`` holds the (processed) flat gradient."""


class GradientWorker(Client):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.__epoch_stat: dict = {}
        self.__momentum_buf: torch.Tensor | None = None
        self.__state: torch.Tensor | None = None
        self.__step = 0
        self.__registered = False

    def _before_training(self) -> None:
        super()._before_training()
        if self.__registered:
            return
        self.__registered = True
        trainer = self.trainer
        assert trainer.hyper_parameter.optimizer_name.lower() == "sgd"
        trainer.append_named_hook(ExecutorHookPoint.OPTIMIZER_STEP, "step", self._step)
        trainer.append_named_hook(ExecutorHookPoint.AFTER_EPOCH, "record", self._record)
        trainer.append_named_hook(ExecutorHookPoint.AFTER_EXECUTE, "report_end", self._report_end)

    def _report_end(self, **kwargs: Any) -> None:
        self.send_data_to_server(Message(end_training=True))

    def _process_gradient(self, gradient: torch.Tensor) -> Any:
        return gradient

    def _decode_gradient(self, payload: Any, numel: int) -> torch.Tensor:
        return payload

    def _apply_fused(self, result: Message, flat: torch.Tensor, momentum_buf: torch.Tensor, lr: float, hp: Any,
                     momentum_ready: bool) -> bool:
        """Override point: decode the server's answer AND update the flat arena in one kernel.  Returns False
        when the subclass has no fused form for this answer (the generic decode + update then runs)."""
        return False

    def _step(self, executor: Any, **kwargs: Any) -> None:
        trainer = executor
        hp = trainer.hyper_parameter
        params = [p for p in trainer.model.parameters()]
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in params]
        real = compute_gradient(params, grads, hp.weight_decay)
        flat = trainer.model_util.flat_parameter
        arena_order = flat is not None and flat.is_cuda and runtime.available()
        if arena_order:
            # the fused update walks the arena, whose 4-D segments are stored channels-last: the
            # exchanged vector must be in that physical order, not the logical (o, c, h, w) one
            layout = trainer.model_util.layout
            names = [n for n, _ in trainer.model.named_parameters()]
            assert names == layout.names
            gradient = torch.empty(layout.total, dtype=torch.float32, device=flat.device)
            for i, g in enumerate(real):
                layout.view_of(gradient, i).copy_(g)
        else:
            gradient = torch.cat([g.reshape(-1) for g in real])
        self.send_data_to_server(GradientMessage(
            in_round=True, other_data={"dataset_size": trainer.dataset_size, "gradient": self._process_gradient(gradient),
                                       "numel": gradient.numel(), "parity": self.__step & 1}))
        result = self._get_data_from_server()
        assert result is not None
        lr = trainer.current_lr()
        if self.__momentum_buf is None or self.__momentum_buf.device != gradient.device:
            self.__momentum_buf = torch.zeros_like(gradient)
            self.__step = 0
        if arena_order and self._apply_fused(result, flat, self.__momentum_buf, lr, hp, self.__step > 0):
            self.__step += 1
            return
        d_p = self._decode_gradient(result.other_data["gradient"], gradient.numel()).to(gradient.device)
        if arena_order:
            dev = flat.device
            if self.__state is None or self.__state.device != dev:
                self.__state = torch.zeros(4, dtype=torch.int32, device=dev)
                self.__offsets = torch.tensor([0, flat.numel()], dtype=torch.int64, device=dev)
            self.__state[1] = 1 if self.__step > 0 else 0
            table = torch.tensor([d_p.data_ptr()], dtype=torch.int64, device=dev)
            lr_t = torch.tensor([lr], dtype=torch.float32, device=dev)
            self.__state[0] = 0
            # weight decay is already inside the exchanged gradient
            sgd_ops.sgd_step(flat, self.__momentum_buf, table, self.__offsets, 1, lr_t, self.__state, 0.0,
                             hp.momentum, 0.0, False)
        else:
            if hp.momentum != 0:
                if self.__step == 0:
                    self.__momentum_buf.copy_(d_p)
                else:
                    self.__momentum_buf.mul_(hp.momentum).add_(d_p)
                d_p = self.__momentum_buf
            with torch.no_grad():
                off = 0
                for p in params:
                    n = p.numel()
                    p.add_(d_p[off: off + n].view_as(p), alpha=-lr)
                    off += n
        self.__step += 1

    def _record(self, **kwargs: Any) -> None:
        epoch = kwargs["epoch"]
        trainer = kwargs["executor"]
        self.__epoch_stat[epoch] = {
            "loss": trainer.performance_metric.get_loss(epoch).data.item(),
            "accuracy": trainer.performance_metric.get_accuracy(epoch).data.item(),
        }
        with open(os.path.join(self.config.save_dir, "epoch_stat.json"), "wt", encoding="utf8") as f:
            json.dump(self.__epoch_stat, f)
