#!/bin/bash
mkdir -p gpurun_out
python tools/knn_debug.py 1000000 0.2 knn_margin=0.1 > gpurun_out/dbg4.log 2>&1
CMD="python bench.py --n 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'knn_group|tile_pass' -s 19 -c 5 -o gpurun_out/r02_prof1 $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
