#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_zz_no_alloc.py -x -q -k "not full_size" > gpurun_out/last_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/last_tests.log
timeout 900 python bench.py --kind clustered --particles 2e6 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/last.json 2> gpurun_out/last.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/last.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), d.get('ms_per_step_median_rank0'), {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','crossing','density','pi','force','sort')})
PY
