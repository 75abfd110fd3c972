"""Flatten one graph-FL session into a one-row table (``exp.txt`` CSV, ``exp.json``, and
``exp.xlsx`` when an Excel writer is installed) -- capability parity with
/root/reference/simulation_lib/analysis/graph_exp_analyzer.py:14-91.

    session_path=<save_dir>/server python -m distributed_learning_simulator_b200.analysis.graph_exp_analyzer
"""
import json
import os
import statistics

import pandas as pd

from .session import GraphSession


def summarize(session: GraphSession) -> dict:
    config = session.config
    res: dict = {
        "exp_name": config.exp_name, "distributed_algorithm": config.distributed_algorithm,
        "dataset_name": config.dc_config.dataset_name, "model_name": config.model_config.model_name,
        "round": config.round, "worker_number": config.worker_number,
    }
    res |= dict(config.algorithm_kwargs or {})
    res |= dict(config.hyper_parameter_config.extra_hyper_parameters or {})
    res["last_test_acc"] = session.last_test_acc
    res["mean_test_acc"] = session.mean_test_acc
    totals: dict = {}
    for data in session.worker_data.values():
        for k, v in data.items():
            if "cnt" not in k and "byte" not in k:
                continue
            if k in ("embedding_bytes", "model_bytes"):
                totals[k] = v
            elif "edge_cnt" in k or "node_cnt" in k:
                totals.setdefault(k, []).append(v)
            else:  # per-round byte counters: sum over workers
                acc = totals.setdefault(k, {})
                for k2, v2 in v.items():
                    acc[k2] = acc.get(k2, 0) + v2
    for k, v in list(totals.items()):
        if isinstance(v, list):
            totals[k] = {"mean": statistics.fmean(v), "std": statistics.stdev(v) if len(v) > 1 else 0.0}
    res |= totals
    res["performance"] = session.round_record
    return {k: json.dumps(v) if isinstance(v, dict) else v for k, v in res.items()}


def write(res: dict, out_dir: str = ".") -> pd.DataFrame:
    df = pd.DataFrame([res])
    txt = os.path.join(out_dir, "exp.txt")
    if os.path.isfile(txt):
        df = pd.concat([pd.read_csv(txt), df], ignore_index=True)
    df.to_csv(txt, index=False)
    df.to_json(os.path.join(out_dir, "exp.json"))
    try:
        df.to_excel(os.path.join(out_dir, "exp.xlsx"), index=False, sheet_name="result")
    except (ImportError, ModuleNotFoundError, ValueError):
        pass  # no Excel writer in this environment
    return df


def main() -> None:
    session_path = os.environ.get("session_path", "").strip()
    assert session_path, "set the environment variable session_path to <save_dir>/server"
    write(summarize(GraphSession(session_path)))


if __name__ == "__main__":
    main()
