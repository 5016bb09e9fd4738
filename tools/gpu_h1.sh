#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_gpu_parity.py -x -q -k "not full_size" > gpurun_out/h1_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/h1_tests.log
timeout 600 python bench.py --particles 1e6 --no-cpu --no-e2e > gpurun_out/h1_bench_1e6.json 2> gpurun_out/h1_bench_1e6.err; echo "bench1e6 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/h1_bench_1e6.json') if l.startswith('{')][0])
print(d['ms_per_step'], {k:v['ms'] for k,v in d['kernels'].items()}, d['per_step'])
PY
