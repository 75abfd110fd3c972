"""``random_label_iid`` sampling: every part draws ``sampled_class_number`` distinct labels at random and the
samples of a label are shared evenly among the parts that drew it (what /root/reference/simulation_lib/sampler/
base.py:9-46 does; python's ``random`` is the generator there as well, so a run seeded through
``config.apply_global_config()`` reproduces its label assignment).

The reference asserts that its single draw happens to cover every label; here the draw is repeated until it
does (a label nobody holds would silently vanish from training).
"""
import random
from typing import Any

from ..engine.dataset import ClassificationDatasetCollection, DatasetCollection, DatasetCollectionSampler
from ..engine.log import get_logger
from ..engine.ml_type import MachineLearningPhase

_MAX_DRAWS = 1000


def draw_label_assignment(labels: list, part_number: int, labels_per_part: int) -> list[list]:
    """``part_number`` label subsets of size ``labels_per_part`` whose union is ``labels``."""
    if not 0 < labels_per_part < len(labels):
        raise ValueError("sampled_class_number must be positive and smaller than the number of labels")
    for _ in range(_MAX_DRAWS):
        assignment = [random.sample(labels, labels_per_part) for _ in range(part_number)]
        if set().union(*assignment) == set(labels):
            return assignment
    raise RuntimeError(f"{part_number} parts of {labels_per_part} labels never covered all {len(labels)} labels")


class RandomLabelIIDSplit(DatasetCollectionSampler):
    def __init__(self, dataset_collection: DatasetCollection, part_number: int, sampled_class_number: int,
                 **kwargs: Any) -> None:
        super().__init__(dataset_collection=dataset_collection, part_number=part_number, **kwargs)
        if not isinstance(dataset_collection, ClassificationDatasetCollection) or dataset_collection.is_mutilabel():
            raise TypeError("random_label_iid needs a single-label classification dataset")
        self.assigned_labels = draw_label_assignment(sorted(dataset_collection.get_labels()), part_number,
                                                     sampled_class_number)
        equal_weights = [{label: 1 for label in part} for part in self.assigned_labels]
        for phase in MachineLearningPhase:
            if phase in self._samplers:
                split = self._samplers[phase].split_indices(parts=equal_weights)
                self._dataset_indices[phase] = dict(enumerate(split))
        training = self._dataset_indices[MachineLearningPhase.Training]
        for worker_id, part in enumerate(self.assigned_labels):
            get_logger().info("worker %s has assigned labels %s", worker_id, part)
            get_logger().info("worker %s has training set size %s", worker_id, len(training[worker_id]))
