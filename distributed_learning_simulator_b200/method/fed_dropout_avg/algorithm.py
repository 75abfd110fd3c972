"""FedDropoutAvg aggregation: every element has its own denominator.

An upload votes for an element only where it kept it (non-zero), weighted by its dataset size; an element that
nobody kept stays zero instead of becoming 0 / 0 (maths of
/root/reference/simulation_lib/method/fed_dropout_avg/algorithm.py:8-19).  On CUDA the same rule is the ``dropout``
variant of the fused reduce kernel (``fused_variant``), which counts the denominators while it sums.
"""
from typing import Any

import torch

from ..common_import import FedAVGAlgorithm


class FedDropoutAvgAlgorithm(FedAVGAlgorithm):
    fused_variant = "dropout"

    def _get_weight(self, dataset_size: int, name: str, parameter: torch.Tensor) -> Any:
        kept = parameter.ne(0)
        return kept.to(torch.float32).mul_(float(dataset_size))

    def _apply_total_weight(self, name: str, parameter: torch.Tensor, total_weight: Any) -> torch.Tensor:
        denominators = total_weight.masked_fill(total_weight.eq(0), 1.0)  # nobody kept it: the sum is 0 as well
        return super()._apply_total_weight(name=name, parameter=parameter, total_weight=denominators)
