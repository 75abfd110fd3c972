"""Flat YAML keys -> nested config objects + factories (``cyy_torch_toolbox.Config``)."""
import copy
import random

import torch
from cyy_naive_lib.log import set_level

from .dataset import ClassificationDatasetCollection
from .executor import HookConfig, HyperParameter, Inferencer, Trainer
from .ml_type import MachineLearningPhase
from .models import create_model


class _NS:
    def __init__(self, **kwargs) -> None:
        self.__dict__.update(kwargs)


_HYPER = ("epoch", "batch_size", "learning_rate", "weight_decay", "momentum", "optimizer_name",
          "learning_rate_scheduler_name", "extra_hyper_parameters")


class Config:
    def __init__(self, **kwargs) -> None:
        self.dc_config = _NS(dataset_name="", dataset_kwargs={})
        self.model_config = _NS(model_name="", model_kwargs={})
        self.hyper_parameter_config = HyperParameter()
        self.trainer_config = _NS(hook_config=HookConfig())
        self.reproducible_env_config = _NS(make_reproducible_env=True, seed=0)
        self.save_dir = ""
        self.log_level = "INFO"
        self.debug = False
        self.cache_transforms = None
        self.save_performance_metric = False
        self.use_slow_performance_metrics = True

    def load_config(self, conf) -> None:
        conf = dict(conf)
        self.dc_config.dataset_name = conf.pop("dataset_name", "")
        self.dc_config.dataset_kwargs = dict(conf.pop("dataset_kwargs", None) or {})
        self.model_config.model_name = conf.pop("model_name", "")
        self.model_config.model_kwargs = dict(conf.pop("model_kwargs", None) or {})
        for k in _HYPER:
            if k in conf:
                setattr(self.hyper_parameter_config, k, conf.pop(k))
        if "use_amp" in conf:
            self.trainer_config.hook_config.use_amp = bool(conf.pop("use_amp"))
        for k, v in conf.items():
            setattr(self, k, dict(v) if isinstance(v, dict) else v)

    def get_save_dir(self) -> str:
        import os

        os.makedirs(self.save_dir, exist_ok=True)
        return self.save_dir

    def apply_global_config(self) -> None:
        set_level(str(self.log_level).upper())
        random.seed(self.reproducible_env_config.seed)
        torch.manual_seed(self.reproducible_env_config.seed)

    def create_dataset_collection(self):
        return ClassificationDatasetCollection(self.dc_config.dataset_name, self.dc_config.dataset_kwargs)

    def _model(self, dc):
        state = torch.random.get_rng_state()
        torch.manual_seed(self.reproducible_env_config.seed + 1)
        try:
            return create_model(self.model_config.model_name, dc.spec)
        finally:
            torch.random.set_rng_state(state)

    def create_trainer(self, dc=None) -> Trainer:
        dc = dc or self.create_dataset_collection()
        return Trainer(self._model(dc), dc, copy.deepcopy(self.hyper_parameter_config),
                       copy.deepcopy(self.trainer_config.hook_config))

    def create_inferencer(self, phase=MachineLearningPhase.Test, dc=None) -> Inferencer:
        dc = dc or self.create_dataset_collection()
        return Inferencer(self._model(dc), dc, phase, copy.deepcopy(self.hyper_parameter_config),
                          copy.deepcopy(self.trainer_config.hook_config))
