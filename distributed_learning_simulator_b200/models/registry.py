from typing import Any, Callable

from torch import nn

_MODELS: dict[str, Callable[..., nn.Module]] = {}


def register_model(name: str, constructor: Callable[..., nn.Module]) -> None:
    _MODELS[name.lower()] = constructor


def get_model_names() -> list[str]:
    return sorted(_MODELS)


def create_model(name: str, dataset_collection: Any, **model_kwargs: Any) -> nn.Module:
    key = name.lower()
    if key not in _MODELS:
        raise KeyError(f"unknown model {name}; known: {get_model_names()}")
    return _MODELS[key](dataset_collection=dataset_collection, **model_kwargs)
