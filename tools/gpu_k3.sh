#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --kind clustered --particles 2e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/k3_plain.log 2>&1; echo "plain rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'list_pass_kernel|knn_kernel' -s 16 -c 5 -o gpurun_out/k3_clustered -f $CMD > gpurun_out/k3_ncu.log 2>&1
echo "ncu rc=$?"
