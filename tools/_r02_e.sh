for cfg in "8 32" "16 32" "8 16" "16 16" "8 64" "4 32"; do
  set -- $cfg
  FLS_E2E_THREADS=$1 FLS_E2E_BATCH=$2 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_$1_$2.json 2>/dev/null
  python - <<PY
import json
d=json.load(open('gpurun_out/e2e_$1_$2.json'))
print("threads $1 batch $2: e2e", d['e2e']['value'], "ms", d['e2e']['ms_per_step'], "flat", d['e2e']['flat']['value'], "resident", d['value'])
PY
done
