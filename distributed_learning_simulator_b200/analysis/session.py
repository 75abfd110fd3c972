"""Read back a finished run from its session directory.

Same contract as /root/reference/simulation_lib/analysis/session.py:9-63: point ``Session`` at
``<save_dir>/server``; it loads ``round_record.json`` and ``config.pkl`` and walks ``..`` for
``worker*`` directories (``hyper_parameter.pk``; ``GraphSession`` also reads
``graph_worker_stat.json``)."""
import functools
import json
import os
import pickle

import dill


class Session:
    def __init__(self, session_dir: str) -> None:
        self.session_dir = session_dir
        with open(os.path.join(session_dir, "round_record.json"), "rt", encoding="utf8") as f:
            self.round_record = {int(k): v for k, v in json.load(f).items()}
        with open(os.path.join(session_dir, "config.pkl"), "rb") as f:
            self.config = pickle.load(f)
        self.worker_data: dict = {}
        for name, path in self._worker_dirs():
            with open(os.path.join(path, "hyper_parameter.pk"), "rb") as f:
                self.worker_data[name] = {"hyper_parameter": dill.load(f)}
        if not self.worker_data:
            raise RuntimeError(f"no worker directories next to {session_dir}")

    def _worker_dirs(self) -> list[tuple[str, str]]:
        parent = os.path.normpath(os.path.join(self.session_dir, ".."))
        return sorted((name, os.path.join(parent, name)) for name in os.listdir(parent)
                      if name.startswith("worker") and os.path.isdir(os.path.join(parent, name)))

    @functools.cached_property
    def rounds(self) -> list:
        return sorted(self.round_record.keys())

    @functools.cached_property
    def last_round(self) -> int:
        return self.rounds[-1]

    @functools.cached_property
    def last_test_acc(self) -> float:
        return self.round_record[self.last_round]["test_accuracy"]

    @functools.cached_property
    def mean_test_acc(self) -> float:
        return sum(r["test_accuracy"] for r in self.round_record.values()) / len(self.round_record)


class GraphSession(Session):
    def __init__(self, session_dir: str) -> None:
        super().__init__(session_dir=session_dir)
        for name, path in self._worker_dirs():
            with open(os.path.join(path, "graph_worker_stat.json"), "rt", encoding="utf8") as f:
                self.worker_data[name] = json.load(f)
