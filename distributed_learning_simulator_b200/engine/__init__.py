"""Training substrate owned by this framework (replaces the external cyy_* libraries the
reference builds on; see SURVEY.md 2.3 for the consumed contract)."""
from .dataset import (ClassificationDatasetCollection, DatasetCollection, DatasetCollectionSampler,  # noqa: F401
                      create_dataset_collection, get_dataset_collection_sampler, global_sampler_factory)
from .device import get_device, get_devices  # noqa: F401
from .executor import HookConfig  # noqa: F401
from .hook import Hook  # noqa: F401
from .hyper_parameter import HyperParameter  # noqa: F401
from .inferencer import Inferencer  # noqa: F401
from .keep_model import KeepModelHook  # noqa: F401
from .log import add_file_handler, get_logger  # noqa: F401
from .ml_type import ExecutorHookPoint, MachineLearningPhase, StopExecutingException  # noqa: F401
from .model_util import ModelUtil, ParameterLayout  # noqa: F401
from .storage import DataStorage, TimeCounter  # noqa: F401
from .tensor import (TensorDict, cat_tensors_to_vector, decompose_tensor_to_list, recursive_tensor_op,  # noqa: F401
                     tensor_to)
from .trainer import Trainer  # noqa: F401
