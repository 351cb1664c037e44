#!/usr/bin/env python
"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X`):
per kernel: launches, total / average duration and share of the first N launches (N = launches of the resident scan
steps; later launches of a bench.py run belong to the host-fed e2e slabs and the fused-aggregate probe).
Usage: python tools/launch_summary.py launches.csv [first_n]"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    first_n = int(sys.argv[2]) if len(sys.argv) > 2 else None
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rows = list(csv.DictReader(lines))
    if first_n:
        rows = rows[:first_n]
    agg = collections.OrderedDict()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = (r["Kernel Name"], r["Grid Size"], r["Block Size"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e3
    total = sum(a[1] for a in agg.values())
    print(f"# {len(rows)} launches, {total:.1f} us in total")
    print("# kernel | launches | total us | avg us | share | grid | block")
    for (name, grid, block), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name} | {n} | {us:.1f} | {us / n:.1f} | {us / total:.3f} | {grid} | {block}")


if __name__ == "__main__":
    main()
