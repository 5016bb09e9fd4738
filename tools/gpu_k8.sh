#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --kind clustered --particles 2e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:'knn_kernel|knn_group_kernel' -s 6 -c 2 -o gpurun_out/k8_knn -f $CMD > gpurun_out/k8_ncu.log 2>&1
echo "ncu rc=$?"
