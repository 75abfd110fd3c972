"""Base experiment configuration: flat YAML keys -> nested config objects + factories.

Own version of ``cyy_torch_toolbox.Config`` (contract reconstructed from
/root/reference/simulation_lib/config.py:8,16-18,34,53,57, practitioner.py:30-31,
server/server.py:34, executor.py:62 and the key inventory in SURVEY.md 5.6).
"""
import os
import random
import threading
from dataclasses import dataclass, field
from typing import Any

import torch

from ..models import create_model
from .dataset import DatasetCollection, create_dataset_collection
from .executor import HookConfig
from .hyper_parameter import HyperParameter
from .inferencer import Inferencer
from .log import set_level
from .ml_type import MachineLearningPhase
from .trainer import Trainer


@dataclass
class DatasetCollectionConfig:
    dataset_name: str = ""
    dataset_kwargs: dict = field(default_factory=dict)


@dataclass
class ModelConfig:
    model_name: str = ""
    model_kwargs: dict = field(default_factory=dict)


@dataclass
class TrainerConfig:
    hook_config: HookConfig = field(default_factory=HookConfig)
    dataloader_kwargs: dict = field(default_factory=dict)


@dataclass
class ReproducibleEnvConfig:
    make_reproducible_env: bool = True
    seed: int = 0


_model_init_lock = threading.Lock()

_HYPER_KEYS = ("epoch", "batch_size", "learning_rate", "weight_decay", "momentum", "optimizer_name",
               "learning_rate_scheduler_name", "extra_hyper_parameters")


class Config:
    def __init__(self, **kwargs: Any) -> None:
        self.dc_config = DatasetCollectionConfig()
        self.model_config = ModelConfig()
        self.hyper_parameter_config = HyperParameter()
        self.trainer_config = TrainerConfig()
        self.reproducible_env_config = ReproducibleEnvConfig()
        self.save_dir: str = ""
        self.log_level: str = "INFO"
        self.debug: bool = False
        self.cache_transforms: str | None = None
        self.save_performance_metric: bool = False
        self.use_slow_performance_metrics: bool = True
        for k, v in kwargs.items():
            setattr(self, k, v)

    # ---- loading -------------------------------------------------------------------
    def load_config(self, conf: dict) -> None:
        conf = dict(conf)
        self.dc_config.dataset_name = conf.pop("dataset_name", self.dc_config.dataset_name)
        self.dc_config.dataset_kwargs = dict(conf.pop("dataset_kwargs", None) or self.dc_config.dataset_kwargs)
        self.model_config.model_name = conf.pop("model_name", self.model_config.model_name)
        self.model_config.model_kwargs = dict(conf.pop("model_kwargs", None) or self.model_config.model_kwargs)
        for k in _HYPER_KEYS:
            if k in conf:
                v = conf.pop(k)
                if k == "extra_hyper_parameters":
                    v = dict(v or {})
                setattr(self.hyper_parameter_config, k, v)
        hook = self.trainer_config.hook_config
        for k in ("use_amp", "use_cuda_graph", "use_fused_kernels"):
            if k in conf:
                setattr(hook, k, bool(conf.pop(k)))
        if "dataloader_kwargs" in conf:
            self.trainer_config.dataloader_kwargs = dict(conf.pop("dataloader_kwargs") or {})
        if "make_reproducible_env" in conf:
            self.reproducible_env_config.make_reproducible_env = bool(conf.pop("make_reproducible_env"))
        if "seed" in conf:
            self.reproducible_env_config.seed = int(conf.pop("seed"))
        for k, v in conf.items():
            if hasattr(self, k):
                cur = getattr(self, k)
                if isinstance(cur, dict) and isinstance(v, dict):
                    v = dict(v)
                setattr(self, k, v)
            else:
                raise KeyError(f"unknown config key {k}")
        hook.debug = bool(self.debug)

    def get_save_dir(self) -> str:
        if not self.save_dir:
            self.save_dir = os.path.join("session", "default")
        os.makedirs(self.save_dir, exist_ok=True)
        return self.save_dir

    def apply_global_config(self) -> None:
        set_level("DEBUG" if self.debug else str(self.log_level).upper())
        if self.reproducible_env_config.make_reproducible_env:
            seed = self.reproducible_env_config.seed
            random.seed(seed)
            torch.manual_seed(seed)
            try:
                import numpy as np

                np.random.seed(seed)
            except ImportError:  # pragma: no cover
                pass

    # ---- factories -------------------------------------------------------------------
    def create_dataset_collection(self) -> DatasetCollection:
        return create_dataset_collection(self.dc_config.dataset_name, self.dc_config.dataset_kwargs)

    def _create_model(self, dc: DatasetCollection) -> torch.nn.Module:
        # every executor of a run starts from the same initial weights; the global torch RNG is
        # process-wide, so concurrent role threads must not interleave here
        with _model_init_lock:
            state = torch.random.get_rng_state()
            torch.manual_seed(self.reproducible_env_config.seed + 1)
            try:
                return create_model(self.model_config.model_name, dc, **self.model_config.model_kwargs)
            finally:
                torch.random.set_rng_state(state)

    def _hook_config(self) -> HookConfig:
        import copy

        return copy.deepcopy(self.trainer_config.hook_config)

    def create_trainer(self, dc: DatasetCollection | None = None) -> Trainer:
        if dc is None:
            dc = self.create_dataset_collection()
        hp = self.hyper_parameter_config.copy()
        if self.model_config.model_kwargs.get("max_len") and dc.spec.get("kind") == "text":
            pass
        trainer = Trainer(self._create_model(dc), dc, hp, self._hook_config())
        trainer.update_dataloader_kwargs(**self.trainer_config.dataloader_kwargs)
        if "num_neighbor" in hp.extra_hyper_parameters:
            trainer.update_dataloader_kwargs(num_neighbor=hp.extra_hyper_parameters["num_neighbor"])
        trainer.set_seed(self.reproducible_env_config.seed)
        return trainer

    def create_inferencer(self, phase: MachineLearningPhase = MachineLearningPhase.Test,
                          dc: DatasetCollection | None = None) -> Inferencer:
        if dc is None:
            dc = self.create_dataset_collection()
        hp = self.hyper_parameter_config.copy()
        hp.batch_size = int(hp.extra_hyper_parameters.get("inference_batch_size", max(hp.batch_size, 1000)))
        inferencer = Inferencer(self._create_model(dc), dc, phase, hp, self._hook_config())
        inferencer.set_seed(self.reproducible_env_config.seed)
        return inferencer
