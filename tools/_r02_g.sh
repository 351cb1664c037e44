python -m pytest tests/test_sql_extension.py -m gpu -x -q > gpurun_out/sql_suite.log 2>&1; echo suite=$?; tail -5 gpurun_out/sql_suite.log; tail -6 gpurun_out/sqltest_extension.log
python tools/lineitem_dbgen.py --sf 10 --parity-sf 1 --sql-runs 5 --out gpurun_out/r02_lineitem_dbgen_sf10.json > gpurun_out/dbgen_sf10.log 2>&1; echo dbgen=$?; tail -3 gpurun_out/dbgen_sf10.log
