#!/bin/bash
mkdir -p gpurun_out
for cfg in "--kind clustered --particles 2e6 --steps 10 --warmup 3"; do
timeout 900 python bench.py $cfg --no-cpu --no-e2e > gpurun_out/k5.json 2> gpurun_out/k5.err; echo "$cfg rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/k5.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items() if k in ('knn','crossing','density','pi','force','sort')}, d['per_step'])
PY
done
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:'knn_kernel|knn_group_kernel' -s 6 -c 2 --csv --log-file gpurun_out/k5_knn.csv python bench.py --kind clustered --particles 2e6 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/k5_ncu.log 2>&1; echo "ncu rc=$?"; grep -v "^==" gpurun_out/k5_knn.csv | cut -d, -f5,13- | tail -4
