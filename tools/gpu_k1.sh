#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --kind clustered --particles 4e6 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/k1_cl4e6.json 2> gpurun_out/k1_cl4e6.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/k1_cl4e6.json') if l.startswith('{')][0])
print(d['ms_per_step'], {k:v['ms'] for k,v in d['kernels'].items()}, d['per_step'])
PY
