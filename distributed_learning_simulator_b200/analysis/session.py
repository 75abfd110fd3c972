"""Reading a finished run back from disk.

A run leaves ``<save_dir>/server/{round_record.json, config.pkl}`` and one ``<save_dir>/worker_<i>/`` per worker
(``hyper_parameter.pk``; graph workers also ``graph_worker_stat.json``).  ``Session(<save_dir>/server)`` and
``GraphSession`` expose them the way /root/reference/simulation_lib/analysis/session.py:9-63 does
(``round_record``, ``config``, ``worker_data``, ``rounds``, ``last_round``, ``last_test_acc``,
``mean_test_acc``)."""
import json
import os
import pickle
from typing import Any, Callable, Iterator

import dill


def _read(path: str, mode: str, loader: Callable) -> Any:
    with open(path, mode, **({"encoding": "utf8"} if "t" in mode else {})) as f:
        return loader(f)


class Session:
    worker_file = "hyper_parameter.pk"

    def __init__(self, session_dir: str) -> None:
        self.session_dir = session_dir
        records = _read(os.path.join(session_dir, "round_record.json"), "rt", json.load)
        self.round_record = {int(round_key): stat for round_key, stat in records.items()}
        self.config = _read(os.path.join(session_dir, "config.pkl"), "rb", pickle.load)
        self.worker_data: dict = {name: self._load_worker(path) for name, path in self._worker_dirs()}
        if not self.worker_data:
            raise RuntimeError(f"no worker directories next to {session_dir}")

    def _worker_dirs(self) -> Iterator[tuple[str, str]]:
        run_dir = os.path.normpath(os.path.join(self.session_dir, os.pardir))
        for entry in sorted(os.listdir(run_dir)):
            path = os.path.join(run_dir, entry)
            if entry.startswith("worker") and os.path.isdir(path):
                yield entry, path

    def _load_worker(self, path: str) -> dict:
        return {"hyper_parameter": _read(os.path.join(path, self.worker_file), "rb", dill.load)}

    # ---- derived views -------------------------------------------------------------------------------------
    @property
    def rounds(self) -> list:
        return sorted(self.round_record)

    @property
    def last_round(self) -> int:
        return max(self.round_record)

    @property
    def last_test_acc(self) -> float:
        return self.round_record[self.last_round]["test_accuracy"]

    @property
    def mean_test_acc(self) -> float:
        accuracies = [stat["test_accuracy"] for stat in self.round_record.values()]
        return sum(accuracies) / len(accuracies)


class GraphSession(Session):
    worker_file = "graph_worker_stat.json"

    def _load_worker(self, path: str) -> dict:
        return _read(os.path.join(path, self.worker_file), "rt", json.load)
