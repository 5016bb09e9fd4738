#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py -x -q -k "stale or moving or variants" > gpurun_out/k2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/k2_tests.log
for ow in 1 0; do
for cfg in "--kind clustered --particles 4e6 --steps 10 --warmup 3" "--particles 1e7 --steps 10 --warmup 3"; do
SPHB200_OPTIONS=knn_order_worklist=$ow timeout 900 python bench.py $cfg --no-cpu --no-e2e > gpurun_out/k2.json 2> gpurun_out/k2.err; echo "ow=$ow $cfg rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/k2.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','crossing','density','pi','force')})
PY
done; done
