python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/parity.log 2>&1; echo parity=$?; tail -12 gpurun_out/parity.log
python bench.py --steps 20 --no-cpu-baseline > gpurun_out/bench_maps.json 2>gpurun_out/bench_maps.err; echo bench=$?; tail -3 gpurun_out/bench_maps.err
FLSGPU_KIND_MAPS=0 python bench.py --steps 20 --no-cpu-baseline > gpurun_out/bench_nomaps.json 2>/dev/null; echo bench0=$?
python - <<PY
import json
for n in ("maps","nomaps"):
    d=json.load(open(f"gpurun_out/bench_{n}.json"))
    print(n, d["value"], d["ms_per_step"], [(c["column"], c["ms"]) for c in d["per_column"]])
PY
