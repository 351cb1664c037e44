python -m pytest tests/test_gpu_strings.py tests/test_gpu_str_filters.py tests/test_zz_no_alloc.py tests/test_reader.py -m gpu -x -q > gpurun_out/str_suite.log 2>&1; echo suite=$?; tail -4 gpurun_out/str_suite.log
for k in 1 0 1 0; do FLSGPU_STR_FUSE_KEYS=$k python bench.py --steps 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('fuse_keys $k', d['value'], d['ms_per_step'], [c['ms'] for c in d['per_column'] if c['column']=='varchar'], d['gpu_launches'])"; done
