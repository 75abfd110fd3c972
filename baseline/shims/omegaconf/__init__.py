"""Minimal omegaconf stand-in: YAML -> dict with ``merge_with`` (reference call sites:
simulation_lib/config.py:86-87,102)."""
import yaml


class DictConfig(dict):
    def merge_with(self, other) -> None:
        def merge(a, b):
            for k, v in b.items():
                if isinstance(v, dict) and isinstance(a.get(k), dict):
                    merge(a[k], v)
                else:
                    a[k] = v

        merge(self, other)


class OmegaConf:
    @staticmethod
    def load(path) -> DictConfig:
        with open(path, "rt", encoding="utf8") as f:
            return DictConfig(yaml.safe_load(f) or {})

    @staticmethod
    def to_container(conf, resolve: bool = True) -> dict:
        return dict(conf)

    @staticmethod
    def create(obj=None) -> DictConfig:
        return DictConfig(obj or {})
