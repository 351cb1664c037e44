python bench.py --steps 10 --no-cpu-baseline > gpurun_out/exp_v2.json 2> gpurun_out/exp_v2.err
python tools/sweep.py --kinds ALP0,ALPX,ALPRD --types DOUBLE,FLOAT --values 134217728 > gpurun_out/exp_sweep_v2.log 2>&1
