"""A participant identity decoupled from the per-task ``worker_id``.

Same role as /root/reference/simulation_lib/practitioner.py:5-35: holds one dataset
sampler per dataset name and builds a trainer whose collection is restricted to this
practitioner's shard.
"""
import threading
from typing import Any

from .engine.dataset import DatasetCollectionSampler
from .engine.trainer import Trainer


_sampler_lock = threading.Lock()


class Practitioner:
    def __init__(self, practitioner_id: int) -> None:
        self.__id: int = practitioner_id
        self.__worker_id: int = practitioner_id
        self._dataset_sampler: dict[str, DatasetCollectionSampler] = {}

    @property
    def id(self) -> int:
        return self.__id

    @property
    def worker_id(self) -> int:
        return self.__worker_id

    def set_worker_id(self, worker_id: int) -> None:
        self.__worker_id = worker_id

    def set_sampler(self, name: str, sampler: DatasetCollectionSampler) -> None:
        assert name not in self._dataset_sampler
        self._dataset_sampler[name] = sampler

    def has_dataset(self, name: str) -> bool:
        return name in self._dataset_sampler

    def create_trainer(self, config: Any) -> Trainer:
        dc = config.create_dataset_collection()
        trainer = config.create_trainer(dc=dc)
        sampler = self._dataset_sampler.get(trainer.dataset_collection.name)
        if sampler is None:
            sampler = self._dataset_sampler[config.dc_config.dataset_name]
        with _sampler_lock:  # samplers are shared between the practitioners of a task
            sampler.set_dataset_collection(trainer.dataset_collection)
            sampler.sample(part_id=self.__worker_id)
        if getattr(config, "merge_validation_to_training_set", False):
            trainer.dataset_collection.merge_validation_to_training()
        trainer.set_seed(config.reproducible_env_config.seed + 1000 * (self.__worker_id + 1))
        trainer.limited_resource = bool(getattr(config, "limited_resource", False))
        return trainer
