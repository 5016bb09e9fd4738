#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/f4_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/f4_tests.log
timeout 900 python bench.py > gpurun_out/f4_bench_1e7.json 2> gpurun_out/f4_bench_1e7.err; echo "bench1e7 rc=$?"; tail -2 gpurun_out/f4_bench_1e7.err
timeout 600 python bench.py --particles 1e6 --no-cpu > gpurun_out/f4_bench_1e6.json 2> gpurun_out/f4_bench_1e6.err; echo "bench1e6 rc=$?"
timeout 600 python bench.py --kind clustered --particles 4e6 --no-cpu --no-e2e > gpurun_out/f4_bench_cl4e6.json 2> gpurun_out/f4_bench_cl4e6.err; echo "cl4e6 rc=$?"
python - <<'PY'
import json
for f in ('f4_bench_1e7','f4_bench_1e6','f4_bench_cl4e6'):
    d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][0])
    print(f, round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items()}, d.get('e2e',{}).get('ms_per_step'))
PY
