"""Hyper-parameters, optimizer and LR-scheduler factories.

The reference configs only ever use ``optimizer_name: SGD`` with
``learning_rate_scheduler_name: CosineAnnealingLR``; graph configs omit
both and fall back to the same defaults.  The cosine schedule is stepped once per
optimizer step over ``epoch * batches_per_epoch`` steps.  ``lr_table()`` exposes the whole
schedule as a tensor so the CUDA-graph training path can read the step's learning rate
from device memory instead of baking a Python float into the captured graph.
"""
import copy
import math
from dataclasses import dataclass, field
from typing import Any

import torch


@dataclass
class HyperParameter:
    epoch: int = 1
    batch_size: int = 64
    learning_rate: float = 0.01
    weight_decay: float = 0.0
    momentum: float = 0.9
    optimizer_name: str = "SGD"
    learning_rate_scheduler_name: str = "CosineAnnealingLR"
    extra_hyper_parameters: dict = field(default_factory=dict)

    def copy(self) -> "HyperParameter":
        return copy.deepcopy(self)

    def get_optimizer(self, parameters: Any) -> torch.optim.Optimizer:
        name = self.optimizer_name.lower()
        if name == "sgd":
            return torch.optim.SGD(parameters, lr=self.learning_rate, momentum=self.momentum,
                                   weight_decay=self.weight_decay)
        if name == "adam":
            return torch.optim.Adam(parameters, lr=self.learning_rate, weight_decay=self.weight_decay)
        if name == "adamw":
            return torch.optim.AdamW(parameters, lr=self.learning_rate, weight_decay=self.weight_decay)
        raise NotImplementedError(self.optimizer_name)

    def get_lr_scheduler(self, optimizer: torch.optim.Optimizer, total_steps: int) -> Any:
        name = self.learning_rate_scheduler_name.lower()
        total_steps = max(total_steps, 1)
        if name == "cosineannealinglr":
            return torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=total_steps)
        if name in ("none", "constant", "constantlr"):
            return torch.optim.lr_scheduler.LambdaLR(optimizer, lambda _: 1.0)
        raise NotImplementedError(self.learning_rate_scheduler_name)

    def lr_table(self, total_steps: int, base_lr: float | None = None) -> torch.Tensor:
        """``lr[t]`` used by optimizer step ``t`` (0-based), matching ``get_lr_scheduler``."""
        base = self.learning_rate if base_lr is None else base_lr
        total_steps = max(total_steps, 1)
        t = torch.arange(total_steps + 1, dtype=torch.float64)
        if self.learning_rate_scheduler_name.lower() == "cosineannealinglr":
            lr = base * (1.0 + torch.cos(math.pi * t / total_steps)) / 2.0
        else:
            lr = torch.full_like(t, base)
        return lr.to(torch.float32)
