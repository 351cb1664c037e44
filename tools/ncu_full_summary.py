#!/usr/bin/env python
"""Summarise an `ncu --set full` report (read with `ncu -i X.ncu-rep --page raw --csv > raw.csv`): one line per captured launch
with the metrics the roofline discussion uses.  Usage: python tools/ncu_full_summary.py raw.csv [label,label,...]"""
import csv
import sys

WANT = [
    "Kernel Name", "launch__grid_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second",
    "FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main():
    rows = list(csv.reader(open(sys.argv[1], newline="")))
    rows = [r for r in rows if r and not r[0].startswith("==")]
    head, units, data = rows[0], rows[1], rows[2:]
    labels = sys.argv[2].split(",") if len(sys.argv) > 2 else []
    idx = [(w, head.index(w)) for w in WANT if w in head]
    print("# launch | " + " | ".join(f"{w} [{units[i]}]" for w, i in idx))
    for n, r in enumerate(data):
        lab = labels[n] if n < len(labels) else str(n)
        print(lab + " | " + " | ".join(r[i] for _, i in idx))


if __name__ == "__main__":
    main()
