"""Stand-in for the private ``cyy_torch_toolbox`` (plain PyTorch; see ../README.md)."""
from .config import Config  # noqa: F401
from .dataset import ClassificationDatasetCollection, DatasetCollection, DatasetCollectionSampler  # noqa: F401
from .executor import Inferencer, Trainer  # noqa: F401
from .ml_type import ExecutorHookPoint, MachineLearningPhase, StopExecutingException  # noqa: F401
