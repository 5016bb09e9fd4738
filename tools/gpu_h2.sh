#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:'tile_pass_kernel' -s 8 -c 4 -o gpurun_out/r02_tile_ring -f $CMD > gpurun_out/ncu_run3.log 2>&1
echo "ncu rc=$?"
