"""A participant of federated tasks.  Its identity (``id``) outlives the tasks it joins; inside one task it is
worker number ``worker_id`` (roles of /root/reference/simulation_lib/practitioner.py:5-35).

A practitioner owns a *claim* on data, not data: per dataset name a sampler that knows which indices belong to
which part.  ``create_trainer`` turns the claim into a trainer whose dataset collection holds only this
participant's shard, seeded per worker so that co-located workers draw different batches.
"""
import threading
from typing import Any

from .engine.dataset import DatasetCollectionSampler
from .engine.trainer import Trainer

# samplers are shared by all practitioners of a task and re-pointed at each new dataset collection
_SAMPLER_LOCK = threading.Lock()


class Practitioner:
    def __init__(self, practitioner_id: int) -> None:
        self.__identity = practitioner_id
        self.__task_worker_id = practitioner_id     # until a task re-numbers its participants
        self._dataset_sampler: dict[str, DatasetCollectionSampler] = {}

    @property
    def id(self) -> int:
        return self.__identity

    @property
    def worker_id(self) -> int:
        return self.__task_worker_id

    def set_worker_id(self, worker_id: int) -> None:
        self.__task_worker_id = worker_id

    # ---- data claims ---------------------------------------------------------------------------------------
    def set_sampler(self, name: str, sampler: DatasetCollectionSampler) -> None:
        if name in self._dataset_sampler:
            raise ValueError(f"practitioner {self.__identity} already has a sampler for {name}")
        self._dataset_sampler[name] = sampler

    def has_dataset(self, name: str) -> bool:
        return name in self._dataset_sampler

    def __sampler_for(self, *names: str) -> DatasetCollectionSampler:
        for name in names:
            if name in self._dataset_sampler:
                return self._dataset_sampler[name]
        raise KeyError(f"practitioner {self.__identity} holds no data of {names}")

    # ---- trainer -------------------------------------------------------------------------------------------
    def create_trainer(self, config: Any) -> Trainer:
        trainer = config.create_trainer(dc=config.create_dataset_collection())
        collection = trainer.dataset_collection
        sampler = self.__sampler_for(collection.name, config.dc_config.dataset_name)
        with _SAMPLER_LOCK:
            sampler.set_dataset_collection(collection)
            sampler.sample(part_id=self.__task_worker_id)
        if getattr(config, "merge_validation_to_training_set", False):
            collection.merge_validation_to_training()
        trainer.set_seed(config.reproducible_env_config.seed + 1000 * (self.__task_worker_id + 1))
        trainer.limited_resource = bool(getattr(config, "limited_resource", False))
        return trainer
