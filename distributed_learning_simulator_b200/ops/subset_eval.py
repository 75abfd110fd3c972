"""Batched evaluation of coalition models for the Shapley engines.
``make_batch_evaluator`` returns None when no batched path applies; the engines then
evaluate coalitions one by one through ``Server.get_metric``."""
from typing import Any


def make_batch_evaluator(server: Any, all_worker_data: dict, metric_type: str) -> Any:
    return None
