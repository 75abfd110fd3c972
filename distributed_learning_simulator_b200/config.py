"""Federated-training configuration and the hydra-compatible loader.

Feature parity with /root/reference/simulation_lib/config.py:16-104:
``DistributedTrainingConfig`` (FL fields, ``load_config_and_process`` deriving
``save_dir``/``log_file``, ``create_practitioners``), ``global_config``, ``load_config``
(CLI), ``load_config_from_file``.  hydra/omegaconf cannot be installed here, so the CLI
grammar the reference's scripts use (/root/reference/test.sh:2-10) is parsed directly:

    simulator.py --config-name fed_avg/mnist.yaml ++fed_avg.round=1 ++fed_avg.algorithm_kwargs.x=2
    simulator.py --config-name large_scale/fed_obd/cifar10.yaml

``--config-name a/b/c.yaml`` packages the file under ``a.b`` exactly like hydra's
directory packaging (the reference unwraps it at config.py:93-94), which is why
overrides are spelled with that prefix; the prefix is optional here.
``conf/global.yaml`` is merged underneath the chosen file (config.py:78-88).
"""
import datetime
import os
import sys
import uuid
from typing import Any

import yaml

from .engine.config import Config
from .engine.dataset import ClassificationDatasetCollection
from .engine.device import get_devices
from .practitioner import Practitioner
from .sampler import get_dataset_collection_sampler

CONF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "conf")


class DistributedTrainingConfig(Config):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.exp_name: str = ""
        self.distributed_algorithm: str = ""
        self.worker_number: int = 0
        self.parallel_number: int = len(get_devices())
        self.round: int = 0
        self.dataset_sampling: str = "iid"
        self.dataset_sampling_kwargs: dict[str, Any] = {}
        self.distribute_init_parameters: bool = True
        self.merge_validation_to_training_set: bool = False
        self.log_file: str = ""
        self.limited_resource: bool = False
        self.endpoint_kwargs: dict = {}
        self.algorithm_kwargs: dict = {}
        # B200 additions (no reference counterpart): how roles are hosted and timed out
        self.topology: str = "auto"          # auto | thread | pipe | device
        self.round_timeout: float | None = None  # watchdog on round barriers (seconds)

    def load_config_and_process(self, conf: dict) -> None:
        self.load_config(conf)
        task_time = datetime.datetime.now()
        date_time = f"{task_time:%Y-%m-%d_%H_%M_%S}"
        dataset_name = self.dc_config.dataset_kwargs.get("name", self.dc_config.dataset_name)
        sampling = self.dataset_sampling if isinstance(self.dataset_sampling, str) else "_".join(self.dataset_sampling)
        dir_suffix = os.path.join(self.distributed_algorithm, f"{dataset_name}_{sampling}",
                                  self.model_config.model_name, date_time, str(uuid.uuid4()))
        if self.exp_name:
            dir_suffix = os.path.join(self.exp_name, dir_suffix)
        root = os.environ.get("DLS_OUTPUT_ROOT", "")
        self.save_dir = os.path.join(root, "session", dir_suffix)
        self.log_file = os.path.join(root, "log", dir_suffix) + ".log"
        assert self.reproducible_env_config.make_reproducible_env

    def create_practitioners(self) -> set:
        practitioners = set()
        dataset_collection = self.create_dataset_collection()
        assert isinstance(dataset_collection, ClassificationDatasetCollection)
        sampler = get_dataset_collection_sampler(
            name=self.dataset_sampling, dataset_collection=dataset_collection,
            part_number=self.worker_number, **self.dataset_sampling_kwargs)
        for practitioner_id in range(self.worker_number):
            practitioner = Practitioner(practitioner_id=practitioner_id)
            practitioner.set_sampler(name=self.dc_config.dataset_name, sampler=sampler)
            practitioners.add(practitioner)
        assert practitioners
        return practitioners


global_config: DistributedTrainingConfig = DistributedTrainingConfig()


def _deep_merge(base: dict, other: dict) -> dict:
    out = dict(base)
    for k, v in other.items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = _deep_merge(out[k], v)
        else:
            out[k] = v
    return out


def _load_yaml(path: str) -> dict:
    with open(path, "rt", encoding="utf8") as f:
        return yaml.safe_load(f) or {}


def _process(conf: dict) -> DistributedTrainingConfig:
    global global_config
    merged = _deep_merge(_load_yaml(os.path.join(CONF_DIR, "global.yaml")), conf)
    fresh = DistributedTrainingConfig()
    fresh.load_config_and_process(merged)
    # keep the module-level object identity: callers hold ``global_config`` by reference
    global_config.__dict__.clear()
    global_config.__dict__.update(fresh.__dict__)
    return global_config


def _set_nested(conf: dict, dotted: str, value: Any) -> None:
    keys = dotted.split(".")
    cur = conf
    for k in keys[:-1]:
        nxt = cur.get(k)
        if not isinstance(nxt, dict):
            nxt = {}
            cur[k] = nxt
        cur = nxt
    cur[keys[-1]] = value


def parse_overrides(overrides: list[str], package: list[str]) -> dict:
    """``++pkg.key.sub=value`` / ``+key=value`` / ``key=value`` -> nested dict, with
    values typed YAML-style (``True``, ``1``, ``0.9``, ``[1,2]``)."""
    result: dict = {}
    for item in overrides:
        if "=" not in item:
            raise ValueError(f"override {item!r} is not key=value")
        key, raw = item.split("=", 1)
        key = key.lstrip("+~")
        parts = key.split(".")
        if package and parts[: len(package)] == package:
            parts = parts[len(package):]
        value = yaml.safe_load(raw) if raw != "" else ""
        _set_nested(result, ".".join(parts), value)
    return result


def load_config(argv: list[str] | None = None) -> DistributedTrainingConfig:
    """CLI entry: fills and returns ``global_config``."""
    argv = list(sys.argv[1:] if argv is None else argv)
    config_name = None
    config_path = CONF_DIR
    overrides: list[str] = []
    i = 0
    while i < len(argv):
        a = argv[i]
        if a in ("--config-name", "-cn"):
            config_name = argv[i + 1]
            i += 2
        elif a.startswith("--config-name="):
            config_name = a.split("=", 1)[1]
            i += 1
        elif a in ("--config-path", "--config-dir", "-cp", "-cd"):
            config_path = argv[i + 1]
            i += 2
        elif a.startswith("--config-path=") or a.startswith("--config-dir="):
            config_path = a.split("=", 1)[1]
            i += 1
        else:
            overrides.append(a)
            i += 1
    if config_name is None:
        raise SystemExit("usage: simulator.py --config-name <algo>/<dataset>.yaml [++<algo>.<key>=<value> ...]")
    if not config_name.endswith((".yaml", ".yml")):
        config_name += ".yaml"
    path = config_name if os.path.isabs(config_name) else os.path.join(config_path, config_name)
    conf = _load_yaml(path)
    # files written in packaged form ({"fed_avg": {...}}) are unwrapped like the reference does
    while "dataset_name" not in conf and len(conf) == 1 and isinstance(next(iter(conf.values())), dict):
        conf = next(iter(conf.values()))
    package = [p for p in os.path.dirname(os.path.relpath(path, config_path)).split(os.sep) if p and p != "."]
    conf = _deep_merge(conf, parse_overrides(overrides, package))
    return _process(conf)


def load_config_from_file(config_file: None | str = None, overrides: dict | None = None) -> DistributedTrainingConfig:
    assert config_file is not None
    if not os.path.isabs(config_file) and not os.path.isfile(config_file):
        config_file = os.path.join(CONF_DIR, config_file)
    conf = _load_yaml(config_file)
    if overrides:
        conf = _deep_merge(conf, overrides)
    return _process(conf)


def load_config_from_dict(conf: dict) -> DistributedTrainingConfig:
    """Build a *new* config object (does not touch ``global_config``)."""
    merged = _deep_merge(_load_yaml(os.path.join(CONF_DIR, "global.yaml")), conf)
    cfg = DistributedTrainingConfig()
    cfg.load_config_and_process(merged)
    return cfg
