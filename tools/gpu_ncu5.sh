#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'knn_group_kernel' -s 2 -c 1 -o gpurun_out/r02_prof5 $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"
