#!/bin/bash
mkdir -p gpurun_out
for cfg in 9 0; do
SPHB200_NVCC_EXTRA="-DTP_PROFILE" SPHB200_OPTIONS=tile_cfg=$cfg timeout 600 python bench.py --particles 1e6 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/h5_bench_$cfg.log 2>&1; echo "cfg=$cfg rc=$?"
grep TP_PROFILE gpurun_out/h5_bench_$cfg.log | tail -8
done
