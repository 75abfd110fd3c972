"""Dataset samplers (/root/reference/simulation_lib/sampler/__init__.py:1-7): the engine's
registry plus the ``random_label_iid`` split."""
from ..engine.dataset import get_dataset_collection_sampler, global_sampler_factory  # noqa: F401
from .base import RandomLabelIIDSplit

global_sampler_factory.register("random_label_iid", RandomLabelIIDSplit)
