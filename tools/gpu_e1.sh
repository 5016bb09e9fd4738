#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gravity.py -x -q > gpurun_out/e1_grav.log 2>&1; echo "gravity rc=$?"; tail -8 gpurun_out/e1_grav.log
for p in -3 -1.5 1 3; do
  echo "== grid_pow=$p"; SPHB200_OPTIONS="grid_pow=$p" timeout 300 python bench.py --particles 4e6 --kind clustered --steps 5 --warmup 2 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], {k:round(v['ms'],2) for k,v in d['kernels'].items()}, d['per_step'])"
done
