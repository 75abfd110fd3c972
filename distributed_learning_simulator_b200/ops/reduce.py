"""Front-end of the fused FedAvg reduction (csrc/reduce.cu).  Filled in by the CUDA
milestone; until the library is present every probe answers "not fusable" and the
algorithms take their torch paths."""
from typing import Any

from . import runtime


def can_fuse(worker_data: Any, old_parameter_dict: Any) -> bool:
    return False


def materialize(worker_data: Any) -> Any:
    return worker_data


def fused_aggregate(pending: dict, old_parameter_dict: Any, variant: str) -> Any:
    raise RuntimeError("fused aggregation is not available")


def try_weighted_avg(data_dict: dict, weight_dict: dict) -> Any:
    return None
