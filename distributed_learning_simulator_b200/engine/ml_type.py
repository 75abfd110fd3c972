"""Core enums and control-flow exceptions of the training substrate.

The reference imports these names from the external ``cyy_torch_toolbox.ml_type``
(see /root/reference/simulation_lib/worker/aggregation_worker.py:6-7 and
worker/worker.py:8); this framework owns them.
"""
from enum import Enum, auto


class MachineLearningPhase(str, Enum):
    Training = "training"
    Validation = "validation"
    Test = "test"

    def __str__(self) -> str:  # log lines use the bare phase name
        return self.value


class ExecutorHookPoint(Enum):
    BEFORE_EXECUTE = auto()
    AFTER_EXECUTE = auto()
    BEFORE_EPOCH = auto()
    AFTER_EPOCH = auto()
    BEFORE_BATCH = auto()
    AFTER_FORWARD = auto()
    AFTER_BATCH = auto()
    OPTIMIZER_STEP = auto()
    AFTER_VALIDATION = auto()
    MODEL_LOADED = auto()


class StopExecutingException(Exception):
    """Raised inside a hook to leave ``Trainer.train()`` cleanly."""


class ModelType(Enum):
    Classification = auto()
    GraphNodeClassification = auto()
    TextClassification = auto()
