#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:'knn_group_kernel|tile_pass_kernel|reorder_kernel|integrate_kernel|knn_kernel|unit_geometry_kernel|pack_kernel' -s 14 -c 10 -o gpurun_out/r02_full_step -f $CMD > gpurun_out/ncu_run2.log 2>&1
echo "ncu full rc=$?"
