#!/bin/bash
mkdir -p gpurun_out
for cfg in 9 0; do
SPHB200_OPTIONS=tile_cfg=$cfg ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:'tile_pass_kernel|unit_geometry|list_pass|repack' -s 40 -c 24 --csv --log-file gpurun_out/h4_launch_$cfg.csv python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/h4_run.log 2>&1; echo "cfg=$cfg rc=$?"
done
