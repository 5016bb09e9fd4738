#!/bin/bash
mkdir -p gpurun_out
for cfg in 0 7; do
SPHB200_OPTIONS=tile_cfg=$cfg timeout 600 python bench.py --particles 1e6 --no-cpu --no-e2e > gpurun_out/h3_bench.json 2> gpurun_out/h3_bench.err; echo "cfg=$cfg rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/h3_bench.json') if l.startswith('{')][0])
print(d['ms_per_step'], {k:v['ms'] for k,v in d['kernels'].items() if k in ('crossing','density','pi','force')})
PY
done
