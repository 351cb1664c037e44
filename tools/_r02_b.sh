set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_strings.py tests/test_gpu_filters.py -m gpu -x -q > gpurun_out/parity_desc.log 2>&1; echo parity=$?; tail -2 gpurun_out/parity_desc.log
tools/membw 800 > gpurun_out/r02_membw.txt 2>&1; echo membw=$?
python tools/lineitem_dbgen.py --sf 1 --parity-sf 0.1 --out gpurun_out/r02_lineitem_dbgen_sf1.json > gpurun_out/dbgen_sf1.log 2>&1; echo dbgen1=$?; tail -5 gpurun_out/dbgen_sf1.log
python tools/lineitem_dist.py --sf 100 --out gpurun_out/r02_lineitem_sf100_n1.json > gpurun_out/sf100_n1.log 2>&1; echo n1=$?; tail -3 gpurun_out/sf100_n1.log
