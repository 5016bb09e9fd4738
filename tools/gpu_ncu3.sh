#!/bin/bash
mkdir -p gpurun_out
export SPHB200_OPTIONS="knn_margin=0.1"
CMD="python bench.py --n 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 120 --csv --log-file gpurun_out/r02_launches3.csv $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'tile_pass_kernel.*ForcePass|tile_pass_kernel.*Density' -s 8 -c 2 -o gpurun_out/r02_prof3 $CMD > gpurun_out/ncu_run2.log 2>&1
echo "ncu rc=$?"
