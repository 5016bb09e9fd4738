#!/bin/bash
mkdir -p gpurun_out
BENCH_VERBOSE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --scaling strong --particles 2.5e6 --steps 6 --warmup 3 --no-e2e --no-verify > gpurun_out/s2.json 2> gpurun_out/s2.err; echo "rc=$?"
grep "^step" gpurun_out/s2.err | tail -4
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/s2.json') if l.startswith('{')][0])
print(round(d['ms_per_step'],3), {k:v['ms'] for k,v in d['kernels'].items()}, d['per_step'])
PY
