python bench.py --steps 10 --no-cpu-baseline > gpurun_out/exp_agg.json 2> gpurun_out/exp_agg.err
