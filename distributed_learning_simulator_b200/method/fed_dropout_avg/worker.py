"""Worker side of FedDropoutAvg: the full model goes up with a Bernoulli(1 - dropout_rate) element mask applied.

Protocol and the two scrape-able log lines (``total_num`` / ``send_num``, read by analysis/analyze_log.py) follow
/root/reference/simulation_lib/method/fed_dropout_avg/worker.py:10-30.  On the flat row the mask, the product and
the count of surviving elements are one kernel pass (``ops.misc.dropout_mask``) rather than a Bernoulli tensor, a
multiply and a ``count_nonzero().item()`` per parameter tensor.
"""
from typing import Any

import torch

from ...engine.log import get_logger
from ...engine.model_util import FlatParameter, as_flat
from ...ops import misc as misc_ops
from ..common_import import AggregationWorker, ParameterMessage


def _drop_elements(parameter: Any, keep_probability: float) -> tuple[Any, int, int]:
    """``(masked parameter, elements in total, elements kept)``."""
    flat = as_flat(parameter)
    if flat is not None:
        masked, kept = misc_ops.dropout_mask(flat.flat, keep_probability)
        return FlatParameter(masked, flat.layout), flat.flat.numel(), int(kept.item())
    masked_dict, total, kept_total = {}, 0, 0
    for name, tensor in parameter.items():
        survivors = torch.bernoulli(torch.full_like(tensor, keep_probability))
        masked_dict[name] = tensor * survivors
        total += tensor.numel()
        kept_total += int(torch.count_nonzero(masked_dict[name]).item())
    return masked_dict, total, kept_total


class FedDropoutAvgWorker(AggregationWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._send_parameter_diff = False  # the server needs absolute values: zero means "dropped"
        self.__keep_probability = 1.0 - float(self.config.algorithm_kwargs["dropout_rate"])
        get_logger().warning("use dropout_rate %s", 1.0 - self.__keep_probability)

    def _get_sent_data(self) -> ParameterMessage:
        self._send_parameter_diff = False
        upload = super()._get_sent_data()
        if not isinstance(upload, ParameterMessage):
            raise TypeError("FedDropoutAvg uploads full models")
        upload.parameter, total, kept = _drop_elements(upload.parameter, self.__keep_probability)
        get_logger().warning("total_num %s", total)
        get_logger().warning("send_num %s", kept)
        return upload
