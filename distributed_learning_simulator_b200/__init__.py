"""B200-native single-host simulator for distributed / federated learning.

Public surface (parity with the reference package ``distributed_learning_simulator``,
/root/reference/simulation_lib/__init__.py:1-4 and SURVEY.md appendix B)."""
from .config import (DistributedTrainingConfig, global_config, load_config, load_config_from_dict,  # noqa: F401
                     load_config_from_file)
from .method.algorithm_factory import CentralizedAlgorithmFactory  # noqa: F401
from .practitioner import Practitioner  # noqa: F401
from .training import get_training_result, train  # noqa: F401

__version__ = "0.1.0"
