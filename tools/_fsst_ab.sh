python -m pytest tests/test_gpu_strings.py tests/test_gpu_str_filters.py tests/test_gpu_lineitem.py -m gpu -x -q > gpurun_out/fsst_suite.log 2>&1; echo suite=$?; tail -4 gpurun_out/fsst_suite.log
export LD_LIBRARY_PATH=$PWD/duckdb-fastlanes_b200/extension/bin:$PWD/duckdb-fastlanes_b200
duckdb-fastlanes_b200/extension/bin/duckdb -c "CALL dbgen(sf=3); COPY (SELECT l_comment FROM lineitem) TO '/dev/shm/c.fls' (FORMAT fls);" > /dev/null
for r in 1 2; do for b in 0 1; do FLSGPU_STR_BALANCE=$b python tools/fsst_file_bench.py /dev/shm/c.fls 20; done; done
for b in 0 1; do FLSGPU_STR_BALANCE=$b python tools/str_bench.py --rows 20000000 --cases fsst 2>/dev/null | tail -1 | cut -c1-300; done
rm -f /dev/shm/c.fls
