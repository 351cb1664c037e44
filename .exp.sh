python bench.py --steps 10 --no-cpu-baseline > gpurun_out/exp_fam3.json 2> gpurun_out/exp_fam3.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fls_decode_kernel -s 8 -c 1 -o gpurun_out/prof_alp -f $CMD > gpurun_out/ncu_alp.log 2>&1
