set -x
python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo bench=$?
tail -c 3000 gpurun_out/r02_bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_ref.err; echo ref=$?
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fls_ -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo list=$?
ncu --set full --clock-control none --import-source on -k regex:fls_ -s 42 -c 14 -o gpurun_out/r02_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo full=$?
ls -la gpurun_out
