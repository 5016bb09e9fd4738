#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:'knn_mma_kernel' -s 3 -c 1 -o gpurun_out/r02_knn_mma -f $CMD > gpurun_out/ncu_run4.log 2>&1
echo "ncu rc=$?"
