#!/usr/bin/env python3
"""Turns the raw ncu outputs brought back in gpurun_out/ into the small text summaries that
are committed under profiles/.

  launches  <launches.csv> <out.md>     per-kernel launch count / total time / share of step
                                        (from `ncu --metrics gpu__time_duration.sum --csv`)
  full      <prof.ncu-rep> <out.md>     per-launch DRAM bytes, DRAM %, tensor %, registers,
                                        occupancy from an `ncu --set full` capture

Runs here (no GPU needed): `ncu -i` reads the report.
"""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("nvllm::", "")
    return name


def launches(src: str, dst: str) -> None:
    agg = defaultdict(lambda: [0, 0.0])
    with open(src, newline="") as f:
        rows = [r for r in f if r.startswith('"')]
    rd = csv.DictReader(rows)
    total = 0.0
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r.get("Metric Unit") in ("us", "usecond"):
            ns *= 1e3
        elif r.get("Metric Unit") in ("ms", "msecond"):
            ns *= 1e6
        k = short(r["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += ns
        total += ns
    with open(dst, "w") as o:
        o.write(f"# ncu launch list summary of `{src}`\n\n")
        o.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES with\n"
                "bench.py's `kernels` table, not absolutes.\n\n")
        o.write(f"total {total / 1e6:.3f} ms over {sum(v[0] for v in agg.values())} launches\n\n")
        o.write("| kernel | launches | total ms | avg us | share |\n|---|---:|---:|---:|---:|\n")
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            o.write(f"| `{k}` | {n} | {ns / 1e6:.3f} | {ns / n / 1e3:.2f} | {100 * ns / total:.1f}% |\n")


WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "smsp__cycles_active.avg",
]


def full(src: str, dst: str) -> None:
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader([l for l in raw.splitlines() if l.startswith('"')]))
    head, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    cols = [c for c in WANT if c in idx]
    # tensor-pipe metric names differ between ncu versions: take whatever matches
    cols += [h for h in head if "pipe_tensor" in h and h not in cols and "pct" in h][:3]
    with open(dst, "w") as o:
        o.write(f"# ncu --set full summary of `{src}`\n\n")
        o.write("| # | kernel | grid | " + " | ".join(f"{c} [{units[idx[c]]}]" for c in cols) + " |\n")
        o.write("|---|---|---|" + "---:|" * len(cols) + "\n")
        for n, r in enumerate(data):
            o.write(f"| {n} | `{short(r[idx['Kernel Name']])}` | {r[idx['Grid Size']]} | " +
                    " | ".join(r[idx[c]] for c in cols) + " |\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
