"""Worker side of FedDropoutAvg: full parameters, Bernoulli(1 - dropout_rate) element mask,
``send_num`` / ``total_num`` log lines (scraped by analysis/analyze_log.py; reference:
method/fed_dropout_avg/worker.py:10-30).  The mask, the product and the non-zero count
are one kernel pass over the flat row (``ops.misc.dropout_mask``) instead of a Bernoulli
tensor + multiply + ``count_nonzero().item()`` per parameter tensor."""
from typing import Any

import torch

from ...engine.log import get_logger
from ...engine.model_util import FlatParameter, as_flat
from ...ops import misc as misc_ops
from ..common_import import AggregationWorker, ParameterMessage


class FedDropoutAvgWorker(AggregationWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.__dropout_rate: float = self.config.algorithm_kwargs["dropout_rate"]
        get_logger().warning("use dropout_rate %s", self.__dropout_rate)

    def _get_sent_data(self) -> ParameterMessage:
        self._send_parameter_diff = False
        sent_data = super()._get_sent_data()
        assert isinstance(sent_data, ParameterMessage)
        keep = 1.0 - self.__dropout_rate
        flat = as_flat(sent_data.parameter)
        if flat is not None:
            masked, count = misc_ops.dropout_mask(flat.flat, keep)
            sent_data.parameter = FlatParameter(masked, flat.layout)
            total_num, send_num = flat.flat.numel(), int(count.item())
        else:
            total_num = send_num = 0
            for k, v in sent_data.parameter.items():
                sent_data.parameter[k] = v * torch.bernoulli(torch.full_like(v, keep))
                total_num += v.numel()
                send_num += int(torch.count_nonzero(sent_data.parameter[k]).item())
        get_logger().warning("total_num %s", total_num)
        get_logger().warning("send_num %s", send_num)
        return sent_data
