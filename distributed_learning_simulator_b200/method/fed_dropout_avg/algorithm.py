"""Per-element denominators: an upload weighs in only where it is non-zero, and elements
nobody kept are divided by one (reference: method/fed_dropout_avg/algorithm.py:8-19).
On CUDA the same maths is the ``dropout`` variant of the fused reduce kernel."""
from typing import Any

import torch

from ..common_import import FedAVGAlgorithm


class FedDropoutAvgAlgorithm(FedAVGAlgorithm):
    fused_variant = "dropout"

    def _get_weight(self, dataset_size: int, name: str, parameter: torch.Tensor) -> Any:
        return (parameter != 0).float() * dataset_size

    def _apply_total_weight(self, name: str, parameter: torch.Tensor, total_weight: Any) -> torch.Tensor:
        # avoid dividing by zero, in such case we set weight to 1
        total_weight = torch.where(total_weight == 0, torch.ones_like(total_weight), total_weight)
        return super()._apply_total_weight(name=name, parameter=parameter, total_weight=total_weight)
