#!/bin/bash
mkdir -p gpurun_out
export SPHB200_OPTIONS="knn_margin=0.1"
CMD="python bench.py --n 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'knn_group' -s 4 -c 1 -o gpurun_out/r02_prof2 $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"
