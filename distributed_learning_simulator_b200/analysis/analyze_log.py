"""Scrape run logs: final test accuracy, per-worker training accuracy, message counts and MB
moved per algorithm.  Capability parity with /root/reference/simulation_lib/analysis/
analyze_log.py:14-290; the log-line shapes it keys on are listed in SURVEY.md 5.5 and are
emitted by engine/inferencer.py, engine/trainer.py, method/fed_dropout_avg/worker.py,
method/smafd/worker.py and algorithm/quantization/deterministic.py.

    logfiles="a.log b.log" python -m distributed_learning_simulator_b200.analysis.analyze_log \
        --config-name fed_obd/cifar10.yaml
"""
import os
import re
import statistics
from typing import Any

_PERCENT = re.compile(r"([0-9.]+)%")
_TAIL_NUMBER = re.compile(r"([0-9.]+(?:e[-+]?[0-9]+)?)\s*$")


def _read(path: str) -> list[str]:
    with open(path, "rt", encoding="utf8") as f:
        return f.readlines()


def _mean_std(values: list[float]) -> dict:
    return {"mean": round(statistics.fmean(values), 2),
            "std": round(statistics.stdev(values), 2) if len(values) > 1 else 0.0}


def final_test_accuracy(lines: list[str], algorithm: str) -> float | None:
    for line in reversed(lines):
        hit = "test loss" in line if algorithm == "sign_SGD" else ("test in" in line and "accuracy" in line)
        if hit:
            found = _PERCENT.findall(line)
            if found:
                return float(found[-1])
    return None


def worker_train_accuracy(lines: list[str], worker_number: int) -> dict[int, float]:
    result: dict[int, float] = {}
    for worker_id in range(worker_number):
        pattern = re.compile(rf"worker {worker_id}\b.*train.*accuracy")
        for line in reversed(lines):
            if pattern.search(line):
                found = _PERCENT.findall(line)
                if found:
                    result[worker_id] = float(found[-1])
                    break
    return result


def compute_acc(paths: list[str], config: Any) -> dict:
    finals, per_worker = [], {}
    for path in paths:
        lines = _read(path)
        acc = final_test_accuracy(lines, config.distributed_algorithm)
        assert acc is not None, f"no test accuracy line in {path}"
        finals.append(acc)
        for w, a in worker_train_accuracy(lines, config.worker_number).items():
            per_worker.setdefault(w, []).append(a)
    summary = _mean_std(finals)
    print("test acc", summary["mean"], summary["std"])
    return {"test_acc": summary, "worker_train_acc": per_worker}


def _tail_numbers(lines: list[str], marker: str) -> list[float]:
    out = []
    for line in lines:
        if marker in line:
            m = _TAIL_NUMBER.search(line.strip())
            if m:
                out.append(float(m.group(1)))
    return out


def compute_data_amount(paths: list[str], config: Any, parameter_number: int | None = None,
                        element_size: int = 4) -> dict:
    """Messages and MiB moved by a run, reconstructed from the config and the logged
    compression ratios / ``send_num`` lines."""
    if parameter_number is None:
        trainer = config.create_trainer()
        parameter_number = int(trainer.model_util.get_parameter_list().numel())
    kw = config.algorithm_kwargs
    chosen = kw.get("random_client_number", config.worker_number)
    uploads = config.round * chosen
    downloads = uploads
    init_msgs = config.worker_number
    msg_num = uploads + downloads + init_msgs
    model_mib = parameter_number * element_size / (1024 * 1024)
    algorithm = config.distributed_algorithm.lower()
    amount: Any = 0.0
    if algorithm == "fed_avg":
        amount = model_mib * msg_num
    elif algorithm == "fed_paq":
        msg_num = config.round * chosen * 2 + config.worker_number
        # 1 byte per element up (8-bit levels), full precision down and for the initial model
        amount = (uploads * parameter_number * 1 + (downloads + init_msgs) * parameter_number * element_size) / 2 ** 20
    elif algorithm == "fed_obd_sq":
        second = kw["second_phase_epoch"] * config.worker_number * 2
        msg_num += second
        amount = (uploads * parameter_number * (1 - kw["dropout_rate"]) + downloads * parameter_number
                  + second * parameter_number + init_msgs * parameter_number * element_size) / 2 ** 20
    elif algorithm == "fed_obd":
        msg_num += kw["second_phase_epoch"] * config.worker_number * 2
        per_path = []
        for path in paths:
            lines = _read(path)
            compressed, remaining, rounds_seen, stage_one = 0.0, msg_num, 0, True
            for line in lines:
                if "broadcast NNABQ compression ratio" in line:
                    ratio = float(_TAIL_NUMBER.search(line.strip()).group(1))
                    rounds_seen += 1
                    if rounds_seen <= config.round:
                        compressed += ratio * chosen
                        remaining -= chosen
                    else:
                        stage_one = False
                        if remaining > config.worker_number:
                            compressed += ratio * config.worker_number
                            remaining -= config.worker_number
                elif "worker NNABQ compression ratio" in line:
                    ratio = float(_TAIL_NUMBER.search(line.strip()).group(1))
                    if stage_one:
                        ratio *= 1 - kw["dropout_rate"]
                    compressed += ratio
                    remaining -= 1
            compressed += max(remaining, 0)  # what is left travels uncompressed (initial model)
            per_path.append(model_mib * compressed)
        amount = _mean_std(per_path)
    elif algorithm in ("fed_dropout_avg", "single_model_afd"):
        per_path = []
        for path in paths:
            sent = sum(_tail_numbers(_read(path), "send_num"))
            assert sent > 0, f"no send_num lines in {path}"
            extra = (downloads + init_msgs) * parameter_number if algorithm == "fed_dropout_avg" \
                else init_msgs * parameter_number
            per_path.append((sent + extra) * element_size / 2 ** 20)
        amount = _mean_std(per_path)
    elif algorithm == "sign_sgd":
        amount = None  # per-step traffic: 1 bit up + 1 trit down per parameter per step
    if isinstance(amount, float):
        amount = round(amount, 2)
    return {"msg_num": msg_num, "data_amount": amount}


def main() -> None:
    from ..config import global_config, load_config

    load_config()
    paths = os.environ.get("logfiles", "").split()
    assert paths, "set the environment variable logfiles to a space separated list of log files"
    compute_acc(paths, global_config)
    res = compute_data_amount(paths, global_config)
    print("msg_num is", res["msg_num"])
    print("data_amount is", res["data_amount"])


if __name__ == "__main__":
    main()
