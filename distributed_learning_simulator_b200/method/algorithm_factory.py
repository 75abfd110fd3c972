"""Registry of distributed algorithms.

Same API as /root/reference/simulation_lib/method/algorithm_factory.py:6-79:
``register_algorithm(algorithm_name, client_cls, server_cls, client_endpoint_cls,
server_endpoint_cls, algorithm_cls)``, ``has_algorithm``, ``create_client``,
``create_server`` (which instantiates ``algorithm_cls()`` into the ``algorithm=`` kwarg).
Registration happens as an import side effect of ``method/__init__.py``.
"""
from typing import Any, Callable

from ..topology.endpoint import ClientEndpoint, ServerEndpoint


class CentralizedAlgorithmFactory:
    config: dict[str, dict] = {}

    @classmethod
    def register_algorithm(cls, algorithm_name: str, client_cls: Callable, server_cls: Callable,
                           client_endpoint_cls: None | Callable = None, server_endpoint_cls: None | Callable = None,
                           algorithm_cls: None | Callable = None) -> None:
        assert algorithm_name not in cls.config, f"{algorithm_name} is already registered"
        entry = {
            "client_cls": client_cls,
            "server_cls": server_cls,
            "client_endpoint_cls": client_endpoint_cls or ClientEndpoint,
            "server_endpoint_cls": server_endpoint_cls or ServerEndpoint,
        }
        if algorithm_cls is not None:
            entry["algorithm_cls"] = algorithm_cls
        cls.config[algorithm_name] = entry

    @classmethod
    def has_algorithm(cls, algorithm_name: str) -> bool:
        return algorithm_name in cls.config

    @classmethod
    def get_algorithm_names(cls) -> list[str]:
        return sorted(cls.config)

    @classmethod
    def create_client(cls, algorithm_name: str, kwargs: dict, endpoint_kwargs: dict,
                      extra_kwargs: dict | None = None, extra_endpoint_kwargs: dict | None = None) -> Any:
        config = cls.config[algorithm_name]
        endpoint = config["client_endpoint_cls"](**(endpoint_kwargs | (extra_endpoint_kwargs or {})))
        return config["client_cls"](endpoint=endpoint, **(kwargs | (extra_kwargs or {})))

    @classmethod
    def create_server(cls, algorithm_name: str, kwargs: dict, endpoint_kwargs: dict,
                      extra_kwargs: dict | None = None, extra_endpoint_kwargs: dict | None = None) -> Any:
        config = cls.config[algorithm_name]
        extra_kwargs = dict(extra_kwargs or {})
        endpoint = config["server_endpoint_cls"](**(endpoint_kwargs | (extra_endpoint_kwargs or {})))
        if "algorithm_cls" in config:
            assert "algorithm" not in extra_kwargs
            extra_kwargs["algorithm"] = config["algorithm_cls"]()
        return config["server_cls"](endpoint=endpoint, **(kwargs | extra_kwargs))
