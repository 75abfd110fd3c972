from enum import Enum, auto


class MachineLearningPhase(str, Enum):
    Training = "training"
    Validation = "validation"
    Test = "test"


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


class StopExecutingException(Exception):
    pass
