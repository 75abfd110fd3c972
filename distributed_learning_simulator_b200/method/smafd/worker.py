"""SMAFD worker: every round ONE sub-model (a random set of parameter tensors within the
``(1 - dropout_rate) * P`` budget, the same for all workers because it is drawn from the
round number) is uploaded as a delta; what is withheld is carried in an error-feedback
residual and added to the next round's delta.  Logs ``send_num`` for the traffic scraper
(reference analysis/analyze_log.py:191-209)."""
import random
from typing import Any

import torch

from ...algorithm.random_dropout_algorithm import RandomDropoutAlgorithm
from ...engine.log import get_logger
from ...message import DeltaParameterMessage, ParameterMessageBase
from ...worker.error_feedback_worker import ErrorFeedbackWorker


class SingleModelAFDWorker(ErrorFeedbackWorker):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.__dropout_rate: float = self.config.algorithm_kwargs["dropout_rate"]

    def sparsify(self, sent_data: ParameterMessageBase) -> ParameterMessageBase:
        assert isinstance(sent_data, DeltaParameterMessage)
        delta = dict(sent_data.delta_parameter)
        rng = random.Random(self.config.reproducible_env_config.seed * 100003 + self._round_num)
        algo = RandomDropoutAlgorithm(self.__dropout_rate, rng=rng)
        kept = set(algo.select_names({k: v.numel() for k, v in delta.items()}))
        new_delta: dict[str, torch.Tensor] = {}
        send_num = 0
        for k, v in delta.items():
            total = v + self._error[k] if k in self._error else v
            if k in kept:
                new_delta[k] = total
                self._error.pop(k, None)
                send_num += v.numel()
            else:
                self._error[k] = total.clone() if total is v else total
        sent_data.delta_parameter = new_delta
        get_logger().warning("send_num %s", send_num)
        return sent_data
