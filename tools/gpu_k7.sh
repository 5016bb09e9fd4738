#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_gpu_parity.py -x -q -k "not full_size" > gpurun_out/k7_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/k7_tests.log
for cfg in "--kind clustered --particles 2e6 --steps 10 --warmup 3" "--particles 1e6" "--particles 1e7 --steps 10 --warmup 3"; do
timeout 900 python bench.py $cfg --no-cpu --no-e2e > gpurun_out/k7.json 2> gpurun_out/k7.err; echo "$cfg rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/k7.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','crossing','density','pi','force','sort')}, d['per_step'])
PY
done
