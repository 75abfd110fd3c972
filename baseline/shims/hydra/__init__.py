"""Minimal hydra stand-in: ``@hydra.main`` parses ``--config-name`` and ``++a.b=c`` overrides
and hands the directory-packaged config to the wrapped function (simulation_lib/config.py:91-95)."""
import functools
import inspect
import os
import sys

import yaml


def main(config_path=None, version_base=None, config_name=None):
    def decorator(fn):
        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            if args or kwargs:
                return fn(*args, **kwargs)
            base = os.path.join(os.path.dirname(inspect.getfile(fn)), config_path or ".")
            if not os.path.isdir(base):
                base = os.path.join(os.path.dirname(inspect.getfile(fn)), "conf")
            name, overrides = config_name, []
            argv = sys.argv[1:]
            i = 0
            while i < len(argv):
                if argv[i] == "--config-name":
                    name = argv[i + 1]
                    i += 2
                else:
                    overrides.append(argv[i])
                    i += 1
            with open(os.path.join(base, name), "rt", encoding="utf8") as f:
                conf = yaml.safe_load(f) or {}
            for part in reversed([p for p in os.path.dirname(name).split(os.sep) if p]):
                conf = {part: conf}
            for item in overrides:
                key, raw = item.lstrip("+").split("=", 1)
                cur = conf
                keys = key.split(".")
                for k in keys[:-1]:
                    cur = cur.setdefault(k, {})
                cur[keys[-1]] = yaml.safe_load(raw)
            from omegaconf import DictConfig

            def wrap(d):
                return DictConfig({k: wrap(v) if isinstance(v, dict) else v for k, v in d.items()})

            return fn(wrap(conf))

        return wrapper

    return decorator
