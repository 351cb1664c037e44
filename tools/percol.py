#!/usr/bin/env python
"""Print the per-column table of one or more bench JSON lines (files given as arguments)."""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        print(f, "ERR", e)
        continue
    print(f"== {f}: value {d['value']} GB/s, {d['ms_per_step']} ms/step, e2e {d['e2e']['value']}, roofline {d['roofline']['frac']}")
    print("   " + "  ".join(f"{p['column']}:{p['ms']:.3f}ms/{p['alg_GBps']:.0f}" for p in d["per_column"]))
    if "reduced_scan" in d and "fused_agg_ms_per_gpu" in d["reduced_scan"]:
        print("   fused agg ms", d["reduced_scan"]["fused_agg_ms_per_gpu"])
