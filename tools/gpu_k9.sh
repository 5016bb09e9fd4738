#!/bin/bash
mkdir -p gpurun_out
for opt in "cell_occ=16" "cell_occ=10" "cell_occ=24" "cell_occ=36" "cell_occ=16,grid_pow=-1.5" "cell_occ=16,grid_pow=-6" "cell_occ=24,grid_pow=-1.5"; do
SPHB200_OPTIONS=$opt timeout 900 python bench.py --kind clustered --particles 2e6 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/k9.json 2> gpurun_out/k9.err; echo "$opt rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/k9.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','crossing','density','pi','force','sort')}, d['per_step'])
PY
done
