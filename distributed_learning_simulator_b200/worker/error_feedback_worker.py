"""Worker whose uploads are sparsified with an error-feedback residual.

Scaffold parity with /root/reference/simulation_lib/worker/error_feedback_worker.py:9-19
(requires delta uploads, keeps ``_error``, ``_get_sent_data = sparsify(...)``); the
concrete ``sparsify`` lives in method/smafd.
"""
from typing import Any

from ..engine.tensor import TensorDict
from ..message import ParameterMessageBase
from .aggregation_worker import AggregationWorker


class ErrorFeedbackWorker(AggregationWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        assert self._send_parameter_diff
        self._error: TensorDict = {}

    def _get_sent_data(self) -> ParameterMessageBase:
        return self.sparsify(super()._get_sent_data())

    def sparsify(self, sent_data: ParameterMessageBase) -> ParameterMessageBase:
        raise NotImplementedError()
