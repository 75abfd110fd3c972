#!/usr/bin/env python3
"""Regenerates profiles/sass_evidence.md from `cuobjdump -sass` of the built library: per-kernel counts
of the SASS mnemonics that prove the Blackwell paths (B200_PROFILING.md), plus excerpts of the tensor-core
main loops.  Runs on the CPU box (no GPU needed)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_learning_simulator_b200", "lib", "libdls_b200.so")
COLS = [("UTCHMMA (tcgen05.mma)", r"\bUTCHMMA"), ("LDTM (tcgen05.ld)", r"\bLDTM"), ("UTMALDG (TMA)", r"\bUTMALDG"),
        ("UTCBAR (tcgen05.commit)", r"\bUTCBAR"), ("cluster (UCGABAR / UTCBAR.MULTICAST)", r"UCGABAR|UTCBAR\.MULTICAST"),
        ("SYNCS (mbarrier)", r"\bSYNCS"), ("LDG", r"\bLDG"), ("STG", r"\bSTG"), ("ATOM/RED", r"\b(ATOM|ATOMG|RED|REDG)\b"),
        ("fp64 ops", r"\bD(ADD|MUL|FMA)\b")]


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = {}
    names = re.findall(r"Function : (\S+)", sass)
    if names:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        demangle = dict(zip(names, out))
    kernels: dict = collections.OrderedDict()
    current = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            current = demangle.get(m.group(1), m.group(1))
            current = re.sub(r"\((int|bool|unsigned int)\)", "", current)
            current = re.sub(r"\(.*", "", current)
            kernels[current] = []
            continue
        if current is not None and re.search(r"/\*[0-9a-f]{4}\*/", line):
            kernels[current].append(line)
    rows = []
    for name, lines in kernels.items():
        counts = [sum(1 for l in lines if re.search(rx, l)) for _, rx in COLS]
        rows.append((name, counts, len(lines)))
    rows.sort(key=lambda r: r[0])
    md = ["# SASS evidence (cuobjdump -sass distributed_learning_simulator_b200/lib/libdls_b200.so)", "",
          "Built with `nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo` (CUDA 12.9); regenerate with "
          "`python tools/sass_evidence.py`.",
          "The mnemonics are the ones /opt/skills/guides/B200_PROFILING.md names as proof of the Blackwell paths.", "",
          "| kernel | " + " | ".join(c for c, _ in COLS) + " | instructions |", "|---|" + "---|" * (len(COLS) + 1)]
    for name, counts, n in rows:
        md.append(f"| `{name}` | " + " | ".join(str(c) for c in counts) + f" | {n} |")
    for title, pattern in [("GEMM main loop and epilogue (gemm_bf16_tcgen05_kernel), excerpt", "gemm_bf16_tcgen05_kernel"),
                           ("Coalition GEMM, weight-stationary cluster variant (subset_gemm_stationary_kernel<128, 2>): "
                            "builder stores into the peer CTA's shared memory, multicast commit, tensor-core loop",
                            "subset_gemm_stationary_kernel<128, 2>"),
                           ("Cross-GPU FedAvg reduce (fedavg_reduce_f32_kernel<true>): .SYS-scoped flag traffic on peer memory",
                            "fedavg_reduce_f32_kernel<true>")]:
        hit = next((k for k in kernels if pattern in k), None)
        if hit is None:
            continue
        md += ["", f"## {title}", "", "```"]
        keep = [l.strip()[:140] for l in kernels[hit]
                if re.search(r"UTC|LDTM|UTMALDG|SYNCS\.(PHASECHK|ARRIVE)|UCGABAR|\bST\.E\.128|\.SYS|FENCE|MEMBAR", l)]
        md += keep[:60] + ["```"]
    path = os.path.join(ROOT, "profiles", "sass_evidence.md")
    with open(path, "w", encoding="utf8") as f:
        f.write("\n".join(md) + "\n")
    print(path, len(rows), "kernels")


if __name__ == "__main__":
    main()
