#!/usr/bin/env python3
"""Turns `ncu --page raw --csv` exports and `--metrics gpu__time_duration.sum` launch lists into
the markdown summaries committed under profiles/."""
import collections
import csv
import json
import re
import sys

PEAKS = json.load(open("MEASURED_PEAKS.json"))
HBM = PEAKS["hbm_gbs"]
BF16 = PEAKS["bf16_tflops"]

COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "MB"), ("dram__bytes_write.sum", "MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block")]


def raw_table(path: str) -> str:
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = ["| kernel | " + " | ".join(label for _, label in COLS) + " | achieved GB/s (dram) | of measured copy peak |",
           "|---|" + "---|" * (len(COLS) + 2)]
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])
        vals = []
        d = {}
        for key, _ in COLS:
            if key in idx:
                v = r[idx[key]].replace(",", "")
                u = units[idx[key]]
                try:
                    f = float(v)
                    if u == "ns":
                        f /= 1e3
                    if u == "ms":
                        f *= 1e3
                    if u == "byte":
                        f /= 1e6
                    if u == "Kbyte":
                        f /= 1e3
                    if u == "Gbyte":
                        f *= 1e3
                    d[key] = f
                    vals.append(f"{f:.2f}" if f < 1000 else f"{f:.0f}")
                except ValueError:
                    vals.append(v)
            else:
                vals.append("-")
        us = d.get("gpu__time_duration.sum")
        mb = d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
        gbs = mb / us * 1e3 if us else 0  # MB/us = TB/s -> *1e3 GB/s
        out.append(f"| {name} | " + " | ".join(vals) + f" | {gbs:.0f} | {gbs / HBM:.2f} |")
    return "\n".join(out)


def launch_table(path: str, top: int = 30) -> str:
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1e3 if unit == "ns" else (v * 1e3 if unit == "ms" else v)
        name = re.sub(r"<.*|\(.*", "", row["Kernel Name"])[:80]
        agg[name][0] += 1
        agg[name][1] += v
    total = sum(v[1] for v in agg.values())
    out = [f"total {total:.0f} us over {sum(v[0] for v in agg.values())} launches", "",
           "| share | total us | launches | avg us | kernel |", "|---|---|---|---|---|"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        out.append(f"| {100 * v[1] / total:.1f}% | {v[1]:.0f} | {v[0]} | {v[1] / v[0]:.2f} | `{k}` |")
    return "\n".join(out)


def events_table(path: str) -> str:
    data = json.load(open(path))
    sustained = PEAKS.get("bf16_tflops_sustained", BF16)
    out = [f"| kernel / size | us (device) | us (eager call incl. Python wrapper) | GB/s | of measured copy peak ({HBM} GB/s) | "
           f"TFLOP/s | of measured bf16 peak ({BF16} burst / {sustained} sustained) |", "|---|---|---|---|---|---|---|"]
    for name, e in data.items():
        gbs, tf = e.get("GB/s"), e.get("TFLOP/s")
        out.append(f"| {name} | {e['us']} | {e.get('us_eager', '-')} | {gbs if gbs is not None else '-'} | "
                   f"{(f'{gbs / HBM:.2f}' if gbs is not None else '-')} | {tf if tf is not None else '-'} | "
                   f"{(f'{tf / BF16:.2f} / {tf / sustained:.2f}' if tf is not None else '-')} |")
    return "\n".join(out)


if __name__ == "__main__":
    kind, path = sys.argv[1], sys.argv[2]
    print({"raw": raw_table, "events": events_table}.get(kind, launch_table)(path))
