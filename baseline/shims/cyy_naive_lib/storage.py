import os
import pickle


class DataStorage:
    def __init__(self, data=None, data_path=None) -> None:
        self._data = data
        self._path = data_path
        self._synced = data is None and data_path is not None

    @property
    def data_path(self):
        return self._path

    def set_data_path(self, path) -> None:
        if path != self._path:
            if self._data is None and self._synced and self._path is not None:
                self._data = self._load()
            self._path = path
            self._synced = False

    def set_data(self, data) -> None:
        self._data = data
        self._synced = False

    def mark_new_data(self) -> None:
        self._synced = False

    @property
    def data(self):
        if self._data is None and self._synced and self._path is not None:
            self._data = self._load()
        return self._data

    def _load(self):
        with open(self._path, "rb") as f:
            return pickle.load(f)

    def save(self) -> None:
        if self._data is None or self._path is None or self._synced:
            return
        os.makedirs(os.path.dirname(os.path.abspath(self._path)), exist_ok=True)
        with open(self._path, "wb") as f:
            pickle.dump(self._data, f)
        self._synced = True
