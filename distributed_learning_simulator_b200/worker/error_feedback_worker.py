"""Base of workers that sparsify their delta upload and carry the part they did not send to the next round.

Scaffold of /root/reference/simulation_lib/worker/error_feedback_worker.py:9-19: ``_error`` is the residual store,
``sparsify`` the method a concrete worker supplies (``method/smafd`` here), and every upload passes through it.
Error feedback only makes sense for deltas, hence the constructor check.
"""
from typing import Any

from ..engine.tensor import TensorDict
from ..message import ParameterMessageBase
from .aggregation_worker import AggregationWorker


class ErrorFeedbackWorker(AggregationWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if not self._send_parameter_diff:
            raise ValueError("error feedback works on delta uploads")
        self._error: TensorDict = {}

    def sparsify(self, sent_data: ParameterMessageBase) -> ParameterMessageBase:
        """Returns what actually goes up; the implementation keeps the rest in ``self._error``."""
        raise NotImplementedError(f"{type(self).__name__} must implement sparsify()")

    def _get_sent_data(self) -> ParameterMessageBase:
        dense = super()._get_sent_data()
        return self.sparsify(dense)
