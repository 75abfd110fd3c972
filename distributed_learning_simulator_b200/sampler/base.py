"""``random_label_iid``: every part draws ``sampled_class_number`` random labels; each label's
samples are split evenly between the parts that drew it
(/root/reference/simulation_lib/sampler/base.py:9-46; python ``random`` is the RNG there too,
so a run seeded through ``config.apply_global_config()`` assigns the same labels)."""
import random
from typing import Any

from ..engine.dataset import ClassificationDatasetCollection, DatasetCollection, DatasetCollectionSampler
from ..engine.log import get_logger
from ..engine.ml_type import MachineLearningPhase


class RandomLabelIIDSplit(DatasetCollectionSampler):
    def __init__(self, dataset_collection: DatasetCollection, part_number: int, sampled_class_number: int,
                 **kwargs: Any) -> None:
        super().__init__(dataset_collection=dataset_collection, part_number=part_number, **kwargs)
        assert isinstance(dataset_collection, ClassificationDatasetCollection)
        assert not dataset_collection.is_mutilabel()
        labels = sorted(dataset_collection.get_labels())
        assert sampled_class_number < len(labels)
        for _ in range(1000):
            assigned_labels = [random.sample(labels, sampled_class_number) for _ in range(part_number)]
            if len(labels) == len(set(sum(assigned_labels, start=[]))):
                break
        # Assure that all labels are allocated (the reference asserts on the first draw)
        assert len(labels) == len(set(sum(assigned_labels, start=[]))), "some label was assigned to nobody"
        self.assigned_labels = assigned_labels
        for phase in MachineLearningPhase:
            if phase not in self._samplers:
                continue
            self._dataset_indices[phase] = dict(enumerate(self._samplers[phase].split_indices(
                parts=[{label: 1 for label in part} for part in assigned_labels])))
        for worker_id, part in enumerate(assigned_labels):
            get_logger().info("worker %s has assigned labels %s", worker_id, part)
            get_logger().info("worker %s has training set size %s", worker_id,
                              len(self._dataset_indices[MachineLearningPhase.Training][worker_id]))
