#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/f1_bench_1e7.json 2> gpurun_out/f1_bench_1e7.err; echo "bench1e7 rc=$?"; tail -3 gpurun_out/f1_bench_1e7.err
timeout 600 python bench.py --particles 1e6 --no-cpu > gpurun_out/f1_bench_1e6.json 2> gpurun_out/f1_bench_1e6.err; echo "bench1e6 rc=$?"
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 130 -c 110 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'knn_group_kernel|tile_pass_kernel|reorder_kernel|integrate_kernel|knn_kernel' -s 12 -c 8 -o gpurun_out/r02_full_step $CMD > gpurun_out/ncu_run2.log 2>&1
echo "ncu full rc=$?"
