"""Copy of the last global model kept by servers and workers.

Same API as /root/reference/simulation_lib/util/model_cache.py:10-51
(``cache_parameter_dict``, ``get_parameter_diff``, ``add_parameter_diff``, ``save``,
``get_parameter_path``, ``load_file``) over the lazy ``DataStorage``.  The reference
pulls every tensor to the CPU first (model_cache.py:24,34,42); here the cache stays on
whatever device the model lives on -- with flat payloads ``get_parameter_diff`` is one
subtraction over ``[P]`` in HBM and nothing crosses PCIe.
"""

from ..engine.storage import DataStorage
from ..engine.tensor import TensorDict, tensor_to
from ..message import FlatParameter, as_flat


class ModelCache:
    def __init__(self) -> None:
        self.__parameter_dict: DataStorage = DataStorage()

    @property
    def parameter_dict(self) -> TensorDict:
        return self.__parameter_dict.data

    @property
    def has_data(self) -> bool:
        return self.__parameter_dict.has_data()

    def load_file(self, path: str) -> None:
        self.__parameter_dict = DataStorage(data_path=path)

    def cache_parameter_dict(self, parameter_dict: TensorDict, path: str) -> None:
        self.__parameter_dict.set_data(parameter_dict)
        self.__parameter_dict.set_data_path(path)

    def get_parameter_diff(self, new_parameter_dict: TensorDict) -> TensorDict:
        cached = self.parameter_dict
        new_flat, old_flat = as_flat(new_parameter_dict), as_flat(cached)
        if new_flat is not None and old_flat is not None and new_flat.layout == old_flat.layout \
                and new_flat.flat.device == old_flat.flat.device:
            return FlatParameter(new_flat.flat - old_flat.flat, new_flat.layout)
        return {k: v - cached[k].to(v.device) for k, v in new_parameter_dict.items()}

    def add_parameter_diff(self, parameter_diff: TensorDict, path: str) -> None:
        self.__parameter_dict.save()
        cached = self.parameter_dict
        diff_flat, old_flat = as_flat(parameter_diff), as_flat(cached)
        if diff_flat is not None and old_flat is not None and diff_flat.layout == old_flat.layout:
            new: TensorDict = FlatParameter(old_flat.flat + diff_flat.flat.to(old_flat.flat.device), old_flat.layout)
        else:
            new = {k: v + parameter_diff[k].to(v.device) for k, v in cached.items()}
        self.__parameter_dict.set_data(new)
        self.__parameter_dict.set_data_path(path)

    def save(self) -> None:
        data = self.__parameter_dict.data
        if data is not None and any(t.is_cuda for t in data.values()):
            self.__parameter_dict.set_data(dict(tensor_to(dict(data), device="cpu")))
        self.__parameter_dict.save()

    def get_parameter_path(self) -> str:
        self.save()
        assert self.__parameter_dict.data_path
        return self.__parameter_dict.data_path
