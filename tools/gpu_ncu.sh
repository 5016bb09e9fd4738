#!/bin/bash
# launch list and --set full capture of one step (N = 10^6, uniform, moving) -> profiles/r02_launches.txt, r02_ncu_full_step.txt
mkdir -p gpurun_out
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 135 -c 120 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'knn_group_kernel|tile_pass_kernel|reorder_kernel|integrate_kernel|knn_kernel|unit_geometry_kernel|pack_kernel|active_groups_kernel' -s 14 -c 11 -o gpurun_out/r02_full_step -f $CMD > gpurun_out/ncu_run2.log 2>&1
echo "ncu full rc=$?"
