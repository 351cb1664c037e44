export LD_LIBRARY_PATH=$PWD/duckdb-fastlanes_b200/extension/bin:$PWD/duckdb-fastlanes_b200
duckdb-fastlanes_b200/extension/bin/duckdb -c "CALL dbgen(sf=2); COPY (SELECT l_comment FROM lineitem) TO '/dev/shm/c.fls' (FORMAT fls);" > /dev/null
python tools/fsst_file_bench.py /dev/shm/c.fls 10
ncu --set full --clock-control none --import-source on -k regex:fls_str_heap -s 3 -c 1 -o gpurun_out/r02_fsst python tools/fsst_file_bench.py /dev/shm/c.fls 2 > gpurun_out/fsst_ncu.log 2>&1; echo ncu=$?
rm -f /dev/shm/c.fls
