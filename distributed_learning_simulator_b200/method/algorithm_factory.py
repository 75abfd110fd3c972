"""Registry of distributed algorithms: a name maps to the classes of its two roles, of their endpoints and,
optionally, of the server-side algorithm object.

Public calls are those of /root/reference/simulation_lib/method/algorithm_factory.py:6-79 --
``register_algorithm``, ``has_algorithm``, ``create_client``, ``create_server`` (which builds ``algorithm_cls()``
and hands it to the server as ``algorithm=``).  Methods register themselves when ``method/__init__.py`` imports
their packages.
"""
from dataclasses import dataclass
from typing import Any, Callable

from ..topology.endpoint import ClientEndpoint, ServerEndpoint


@dataclass(frozen=True)
class _Recipe:
    client_cls: Callable
    server_cls: Callable
    client_endpoint_cls: Callable
    server_endpoint_cls: Callable
    algorithm_cls: Callable | None


def _merged(base: dict, extra: dict | None) -> dict:
    return {**base, **(extra or {})}


class CentralizedAlgorithmFactory:
    _recipes: dict[str, _Recipe] = {}

    @classmethod
    def register_algorithm(cls, algorithm_name: str, client_cls: Callable, server_cls: Callable,
                           client_endpoint_cls: None | Callable = None, server_endpoint_cls: None | Callable = None,
                           algorithm_cls: None | Callable = None) -> None:
        if algorithm_name in cls._recipes:
            raise ValueError(f"{algorithm_name} is already registered")
        cls._recipes[algorithm_name] = _Recipe(client_cls, server_cls, client_endpoint_cls or ClientEndpoint,
                                               server_endpoint_cls or ServerEndpoint, algorithm_cls)

    @classmethod
    def has_algorithm(cls, algorithm_name: str) -> bool:
        return algorithm_name in cls._recipes

    @classmethod
    def get_algorithm_names(cls) -> list[str]:
        return sorted(cls._recipes)

    @classmethod
    def __recipe(cls, algorithm_name: str) -> _Recipe:
        try:
            return cls._recipes[algorithm_name]
        except KeyError:
            raise KeyError(f"unknown distributed_algorithm {algorithm_name!r}; registered: "
                           f"{', '.join(cls.get_algorithm_names())}") from None

    @classmethod
    def create_client(cls, algorithm_name: str, kwargs: dict, endpoint_kwargs: dict,
                      extra_kwargs: dict | None = None, extra_endpoint_kwargs: dict | None = None) -> Any:
        recipe = cls.__recipe(algorithm_name)
        endpoint = recipe.client_endpoint_cls(**_merged(endpoint_kwargs, extra_endpoint_kwargs))
        return recipe.client_cls(endpoint=endpoint, **_merged(kwargs, extra_kwargs))

    @classmethod
    def create_server(cls, algorithm_name: str, kwargs: dict, endpoint_kwargs: dict,
                      extra_kwargs: dict | None = None, extra_endpoint_kwargs: dict | None = None) -> Any:
        recipe = cls.__recipe(algorithm_name)
        endpoint = recipe.server_endpoint_cls(**_merged(endpoint_kwargs, extra_endpoint_kwargs))
        role_kwargs = _merged(kwargs, extra_kwargs)
        if recipe.algorithm_cls is not None:
            if "algorithm" in role_kwargs:
                raise ValueError("the registry builds the algorithm object; do not pass one")
            role_kwargs["algorithm"] = recipe.algorithm_cls()
        return recipe.server_cls(endpoint=endpoint, **role_kwargs)
