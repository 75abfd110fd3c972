"""Hook that logs which sub-modules moved between two model loads (debug aid; reference:
/root/reference/simulation_lib/analysis/module_diff.py:8-44).  Attach with
``trainer.append_hook(ModuleDiff(delta=0.1))``: the first snapshot is taken at
``BEFORE_EXECUTE``, every ``MODEL_LOADED`` logs ``module <name> has diff <L2>``."""
from typing import Any

import torch

from ..engine.hook import Hook
from ..engine.log import get_logger
from ..engine.model_util import ModelUtil


class ModuleDiff(Hook):
    def __init__(self, delta: float | None = 0.1) -> None:
        super().__init__()
        self.__prev: list | None = None
        self.__delta = delta
        self.last_report: dict[str, float] = {}

    @staticmethod
    def _snapshot(executor: Any) -> list:
        util = ModelUtil(executor.model)
        return [(name, torch.cat([p.detach().reshape(-1) for p in module.parameters()]).clone())
                for name, module in util.get_sub_modules() if list(module.parameters())]

    def _before_execute(self, executor: Any, **kwargs: Any) -> None:
        if self.__prev is None:
            self.__prev = self._snapshot(executor)

    def _model_loaded(self, executor: Any, **kwargs: Any) -> None:
        cur = self._snapshot(executor)
        if self.__prev is not None:
            assert len(self.__prev) == len(cur)
            self.last_report = {}
            for (prev_name, prev), (name, value) in zip(self.__prev, cur):
                assert prev_name == name
                diff = float(torch.linalg.norm(prev - value))
                if self.__delta is not None and diff <= self.__delta:
                    continue
                self.last_report[name] = diff
                get_logger().info("module %s has diff %s", name, diff)
        self.__prev = cur
