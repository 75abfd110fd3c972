"""Random tensor-level dropout under a parameter budget (used by ``single_model_afd``).

Same intent as /root/reference/simulation_lib/algorithm/random_dropout_algorithm.py:7-31
with defect D5 fixed: the reference never increments ``partial_parameter_num``, so its
budget never binds; here each kept tensor is charged ``numel`` against
``(1 - dropout_rate) * P``.
"""
import random
from typing import Any

from ..engine.log import get_logger


class RandomDropoutAlgorithm:
    def __init__(self, dropout_rate: float, rng: random.Random | None = None) -> None:
        self.__dropout_rate = dropout_rate
        self.__rng = rng or random
        get_logger().warning("use dropout rate %s", self.__dropout_rate)

    def select_names(self, numels: dict[str, int]) -> list[str]:
        parameter_num = sum(numels.values())
        threshold = (1 - self.__dropout_rate) * parameter_num
        partial_parameter_num = 0
        names = list(numels.keys())
        self.__rng.shuffle(names)
        kept = []
        for k in names:
            if partial_parameter_num > threshold:
                break
            if partial_parameter_num + numels[k] > threshold:
                continue
            partial_parameter_num += numels[k]
            kept.append(k)
        get_logger().info("partial_parameter_num %s threshold %s parameter_num %s", partial_parameter_num,
                          threshold, parameter_num)
        return kept

    def drop_parameter(self, parameter_dict: dict[str, Any]) -> dict[str, Any]:
        kept = set(self.select_names({k: v.numel() for k, v in parameter_dict.items()}))
        return {k: v for k, v in parameter_dict.items() if k in kept}
