#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_gpu_parity.py -x -q -k "not full_size" > gpurun_out/h7_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/h7_tests.log
for mma in 1 0; do
SPHB200_OPTIONS=knn_mma=$mma timeout 600 python bench.py --particles 1e6 --no-cpu --no-e2e > gpurun_out/h7_bench.json 2> gpurun_out/h7_bench.err; echo "mma=$mma rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/h7_bench.json') if l.startswith('{')][0])
print(d['ms_per_step'], {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','sort','density')}, d['per_step'])
PY
done
