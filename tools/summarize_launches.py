"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel:  python tools/summarize_launches.py file.csv
Prints a markdown table: launches, total / mean duration, share of the captured GPU time (cold-cache, serialised times)."""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        rows.append((r["Kernel Name"], r["Grid Size"], r["Block Size"], float(r["Metric Value"].replace(",", ""))))
    agg = OrderedDict()
    for name, grid, block, ns in rows:
        short = re.sub(r"^void ", "", name)
        short = re.sub(r"\(.*\)$", "", short)
        key = (short, grid, block)
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print(f"{len(rows)} launches, {total / 1e6:.3f} ms of GPU time captured\n")
    print("| kernel | grid | block | launches | total ms | mean µs | share |")
    print("|---|---|---|---:|---:|---:|---:|")
    for (name, grid, block), (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {grid} | {block} | {n} | {ns / 1e6:.3f} | {ns / n / 1e3:.1f} | {100 * ns / total:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1])
