"""Per-round metric curves across runs and algorithms (capability parity with
/root/reference/simulation_lib/analysis/analyze_round.py:16-70).  The reference collects
``visualizer/**/round_<n>/test/performance_metric.json`` files written by its external
toolbox; this framework's per-round metrics live in ``<save_dir>/server/round_record.json``,
so those are collected instead.  One CSV per metric is always written; a PNG line plot
(mean +- sd over runs, one line per algorithm) is added when matplotlib is importable.

    config_files="fed_avg/cifar10.yaml fed_obd/cifar10.yaml" python -m \
        distributed_learning_simulator_b200.analysis.analyze_round
"""
import json
import os

import pandas as pd


def find_directories(root: str, name: str) -> list[str]:
    """Every directory called ``name`` below ``root`` (``cyy_naive_lib.fs.path.find_directories``)."""
    return sorted(os.path.join(d, name) for d, subdirs, _ in os.walk(root) if name in subdirs)


def extract_data(session_path: str, algorithm: str, frames: dict) -> dict:
    for server_dir in find_directories(session_path, "server"):
        record = os.path.join(server_dir, "round_record.json")
        if not os.path.isfile(record):
            continue
        with open(record, "rt", encoding="utf8") as f:
            rounds = json.load(f)
        for round_no, metrics in rounds.items():
            for metric, value in metrics.items():
                frames.setdefault(metric, []).append({"round": int(round_no), metric: value, "algorithm": algorithm,
                                                      "run": server_dir})
    return frames


def plot(frames: dict, out_dir: str = ".") -> list[str]:
    written = []
    for metric, rows in frames.items():
        df = pd.DataFrame(rows)
        csv = os.path.join(out_dir, f"{metric}.csv")
        df.to_csv(csv, index=False)
        written.append(csv)
        try:
            import matplotlib

            matplotlib.use("Agg")
            import matplotlib.pyplot as plt
        except ImportError:
            continue
        fig, ax = plt.subplots()
        for algorithm, part in df.groupby("algorithm"):
            stat = part.groupby("round")[metric].agg(["mean", "std"]).fillna(0.0)
            ax.plot(stat.index, stat["mean"], label=algorithm)
            ax.fill_between(stat.index, stat["mean"] - stat["std"], stat["mean"] + stat["std"], alpha=0.2)
        ax.set_xlabel("round")
        ax.set_ylabel(metric)
        ax.legend()
        fig.tight_layout()
        png = os.path.join(out_dir, f"{metric}.png")
        fig.savefig(png)
        plt.close(fig)
        written.append(png)
    return written


def main() -> None:
    from ..config import load_config_from_file

    frames: dict = {}
    root = os.environ.get("DLS_OUTPUT_ROOT", "")
    for config_file in os.environ.get("config_files", "").split():
        config = load_config_from_file(config_file=config_file)
        session_path = os.path.join(root, "session", config.distributed_algorithm)
        extract_data(session_path, config.distributed_algorithm, frames)
    for path in plot(frames):
        print("wrote", path)


if __name__ == "__main__":
    main()
