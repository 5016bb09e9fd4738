#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/h6_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/h6_tests.log
timeout 600 python bench.py --particles 1e6 --no-cpu --no-e2e > gpurun_out/h6_bench_1e6.json 2> gpurun_out/h6_bench_1e6.err; echo "bench1e6 rc=$?"
timeout 900 python bench.py --no-cpu > gpurun_out/h6_bench_1e7.json 2> gpurun_out/h6_bench_1e7.err; echo "bench1e7 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/h6_bench_1e6.json','gpurun_out/h6_bench_1e7.json'):
    d=json.loads([l for l in open(f) if l.startswith('{')][0])
    print(d['ms_per_step'], {k:v['ms'] for k,v in d['kernels'].items()}, d.get('e2e',{}).get('ms_per_step'))
PY
