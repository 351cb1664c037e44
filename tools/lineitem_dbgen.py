#!/usr/bin/env python
"""BASELINE config 3 with the REAL table: TPC-H lineitem from DuckDB's dbgen (reference tree duckdb/extension/tpch/dbgen/dbgen.cpp:406-422,
the generator the reference's test/sql/simple.test:17-40 uses), written through the extension's COPY ... (FORMAT fls), read back
through read_fls on the GPU.  Needs the binaries of duckdb-fastlanes_b200/extension/bin (they travel to the GPU box).

  1. duckdb: CALL dbgen(sf); COPY lineitem TO <file> (FORMAT fls)                      -> times, file size
  2. SQL parity (at --parity-sf, default 1): lineitem EXCEPT ALL read_fls(<file>) and the reverse are empty, counts equal
  3. SQL timing, native DuckDB table (CPU, all cores) vs read_fls (GPU): SUM/COUNT of the 8 numeric columns, TPC-H Q1, Q6,
     MAX of all 16 columns (a scan that materialises every column); best of --sql-runs; results compared as text
  4. the resident per-column scans / fused aggregates / Q1-Q6 masks of tools/lineitem_bench.py on that very file
     (answers from DuckDB's own SUMs over the dbgen table)
One JSON document.   python tools/lineitem_dbgen.py --sf 10 --out gpurun_out/r02_lineitem_dbgen_sf10.json"""
import argparse
import csv
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "duckdb-fastlanes_b200", "extension", "bin")
ENV = dict(os.environ, LD_LIBRARY_PATH=BIN + ":" + os.path.join(ROOT, "duckdb-fastlanes_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))

AGG_COLS = ["l_orderkey", "l_partkey", "l_suppkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount", "l_tax"]
Q_COLS = ["l_quantity", "l_extendedprice", "l_discount", "l_tax"]
ALL_COLS = AGG_COLS + ["l_returnflag", "l_linestatus", "l_shipdate", "l_commitdate", "l_receiptdate", "l_shipinstruct", "l_shipmode", "l_comment"]

SQL = {
    "sum_count": "SELECT " + ", ".join("SUM(%s)" % c for c in AGG_COLS) + ", COUNT(*) FROM {src}",
    "q1": "SELECT l_returnflag, l_linestatus, sum(l_quantity), sum(l_extendedprice), sum(l_extendedprice * (1 - l_discount)), "
          "sum(l_extendedprice * (1 - l_discount) * (1 + l_tax)), avg(l_quantity), avg(l_extendedprice), avg(l_discount), count(*) "
          "FROM {src} WHERE l_shipdate <= CAST('1998-09-02' AS date) GROUP BY l_returnflag, l_linestatus ORDER BY l_returnflag, l_linestatus",
    "q6": "SELECT sum(l_extendedprice * l_discount) FROM {src} WHERE l_shipdate >= CAST('1994-01-01' AS date) AND l_shipdate < CAST('1995-01-01' AS date) "
          "AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24",
    "max_all_columns": "SELECT " + ", ".join("MAX(%s)" % c for c in ALL_COLS) + " FROM {src}",
}


def duckdb(db, script, timeout=3000):
    """Run `script` (statements and dot commands, one per line) in the shell on database file `db`; returns stdout."""
    p = subprocess.run([os.path.join(BIN, "duckdb"), db], input=script, capture_output=True, text=True, env=ENV, timeout=timeout)
    if p.returncode != 0 or "Error" in p.stderr:
        raise RuntimeError("duckdb failed:\n" + p.stderr[-2000:] + "\n" + p.stdout[-2000:])
    return p.stdout


def timed_sql(db, stmt, runs, setup=""):
    """best wall time of `runs` executions (the shell's .timer) and the result rows as text (csv mode)."""
    out = duckdb(db, ".mode csv\n" + setup + ".timer on\n" + (stmt + ";\n") * runs)
    times = [float(x) for x in re.findall(r"Run Time \(s\): real ([0-9.]+)", out)]
    rows = [ln for ln in out.splitlines() if ln and not ln.startswith("Run Time")]
    per = len(rows) // runs
    assert len(times) == runs and rows[:per] == rows[-per:], "the runs of one statement disagree"
    return min(times), rows[:per]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=1.0)
    ap.add_argument("--parity-sf", type=float, default=1.0)
    ap.add_argument("--sql-runs", type=int, default=3)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=8.0)
    ap.add_argument("--gpu-devices", type=int, default=1, help="read_fls shards its batches of row groups over this many devices of the one DuckDB process")
    ap.add_argument("--no-resident", action="store_true", help="SQL part only")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    tmp = tempfile.mkdtemp(prefix="fls_dbgen_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    doc = {"workload": "TPC-H lineitem from DuckDB dbgen, SF%g, written by COPY ... (FORMAT fls), read by read_fls" % args.sf}
    try:
        # ---- SQL parity on a table small enough for EXCEPT ALL over all 16 columns
        pdb, pfls = os.path.join(tmp, "parity.duckdb"), os.path.join(tmp, "parity.fls")
        duckdb(pdb, "CALL dbgen(sf=%g);\nCOPY lineitem TO '%s' (FORMAT fls);\n" % (args.parity_sf, pfls))
        out = duckdb(pdb, ".mode csv\n"
                          "SELECT count(*) FROM lineitem;\nSELECT count(*) FROM read_fls('%s');\n"
                          "SELECT count(*) FROM (SELECT * FROM lineitem EXCEPT ALL SELECT * FROM read_fls('%s'));\n"
                          "SELECT count(*) FROM (SELECT * FROM read_fls('%s') EXCEPT ALL SELECT * FROM lineitem);\n" % (pfls, pfls, pfls))
        nums = [int(x) for x in re.findall(r"^(\d+)$", out, re.M)]
        assert len(nums) == 4 and nums[0] == nums[1] and nums[2] == 0 and nums[3] == 0, "read_fls differs from the dbgen table: %s" % nums
        doc["sql_parity"] = {"sf": args.parity_sf, "rows": nums[0], "lineitem_except_all_read_fls": nums[2], "read_fls_except_all_lineitem": nums[3],
                             "note": "all 16 columns, EXCEPT ALL both ways, through the DuckDB extension on the GPU"}
        os.remove(pdb)
        os.remove(pfls)

        # ---- the table proper
        db, fls = os.path.join(tmp, "tpch.duckdb"), os.path.join(tmp, "lineitem.fls")
        t0 = time.perf_counter()
        duckdb(db, "CALL dbgen(sf=%g);\n" % args.sf)
        t1 = time.perf_counter()
        duckdb(db, "COPY lineitem TO '%s' (FORMAT fls);\n" % fls)
        t2 = time.perf_counter()
        doc["dbgen_s"], doc["copy_to_fls_s"] = round(t1 - t0, 1), round(t2 - t1, 1)
        doc["fls_MB"], doc["duckdb_file_MB"] = round(os.path.getsize(fls) / 1e6, 1), round(os.path.getsize(db) / 1e6, 1)

        # ---- SQL timings: the native table on the host cores vs read_fls on the GPU
        setup = "SET fastlanes_gpu_devices='%s';\n" % ",".join(str(g) for g in range(args.gpu_devices)) if args.gpu_devices > 1 else ""
        sql = {}
        for name, stmt in SQL.items():
            cpu_s, cpu_rows = timed_sql(db, stmt.format(src="lineitem"), args.sql_runs)
            gpu_s, gpu_rows = timed_sql(db, stmt.format(src="read_fls('%s')" % fls), args.sql_runs, setup)
            assert cpu_rows[1:] == gpu_rows[1:], "%s: read_fls result differs from the native table:\n%s\n%s" % (name, cpu_rows, gpu_rows)
            sql[name] = {"duckdb_native_table_s": cpu_s, "read_fls_gpu_s": gpu_s, "speedup": round(cpu_s / gpu_s, 2), "result_rows": len(cpu_rows) - 1}
        doc["sql"] = {"host_threads": os.cpu_count(), "gpu_devices": args.gpu_devices, "runs": args.sql_runs, "timing": "the shell's .timer, best run; results equal as text", "queries": sql}

        # ---- answers for the resident bench from DuckDB itself (dates as days since 1970, decimals unscaled)
        want_csv = os.path.join(tmp, "want.csv")
        q1 = "l_shipdate <= DATE '1998-09-02'"
        q6 = "l_shipdate >= DATE '1994-01-01' AND l_shipdate < DATE '1995-01-01' AND l_discount >= 0.05 AND l_discount <= 0.07 AND l_quantity < 24"

        def unscaled(c):
            return "CAST(SUM(%s) * 100 AS HUGEINT)" % c if c in Q_COLS else "CAST(SUM(%s) AS HUGEINT)" % c
        sel = ["COUNT(*) AS n"] + ["%s AS sum_%s" % (unscaled(c), c) for c in AGG_COLS]
        sel += ["COUNT(*) FILTER (WHERE %s) AS q1_count" % q1] + ["CAST(SUM(%s) FILTER (WHERE %s) * 100 AS HUGEINT) AS q1_%s" % (c, q1, c) for c in Q_COLS]
        sel += ["COUNT(*) FILTER (WHERE %s) AS q6_count" % q6] + ["CAST(SUM(%s) FILTER (WHERE %s) * 100 AS HUGEINT) AS q6_%s" % (c, q6, c) for c in Q_COLS]
        duckdb(db, "COPY (SELECT %s FROM lineitem) TO '%s' (HEADER);\n" % (", ".join(sel), want_csv))
        row = next(csv.DictReader(open(want_csv)))
        want = {"sum": {c: int(row["sum_" + c]) for c in AGG_COLS}, "q1": {c: int(row["q1_" + c]) for c in Q_COLS},
                "q6": {c: int(row["q6_" + c]) for c in Q_COLS}, "q1_count": int(row["q1_count"]), "q6_count": int(row["q6_count"])}
        n_rows = int(row["n"])
        os.remove(db)

        if args.no_resident:
            os.remove(fls)
            print(json.dumps(doc, indent=1))
            if args.out:
                with open(args.out, "w") as f:
                    json.dump(doc, f, indent=1)
            return
        # ---- resident scans of the same file (tools/lineitem_bench.py)
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import lineitem_bench
        import numpy as np
        img = np.fromfile(fls, dtype=np.uint8)
        os.remove(fls)
        doc["resident"] = lineitem_bench.measure(img, want, n_rows, args.iters, args.cpu_seconds)
    finally:
        subprocess.call(["rm", "-rf", tmp])
    print(json.dumps(doc, indent=1))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(doc, f, indent=1)


if __name__ == "__main__":
    main()
