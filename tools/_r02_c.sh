set -x
python -m pytest tests/test_gpu_filters.py tests/test_sql_extension.py tests/test_multi_device.py -m gpu -x -q > gpurun_out/sql_suite.log 2>&1; echo sqlsuite=$?; tail -30 gpurun_out/sql_suite.log
tail -40 gpurun_out/sqltest_extension.log
python tools/lineitem_dbgen.py --sf 1 --parity-sf 0.1 --out gpurun_out/r02_lineitem_dbgen_sf1.json > gpurun_out/dbgen_sf1.log 2>&1; echo dbgen1=$?; tail -5 gpurun_out/dbgen_sf1.log
