#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/f2_tests.log
timeout 900 python bench.py > gpurun_out/f2_bench_1e7.json 2> gpurun_out/f2_bench_1e7.err; echo "bench1e7 rc=$?"; tail -3 gpurun_out/f2_bench_1e7.err
timeout 600 python bench.py --particles 1e6 --no-cpu > gpurun_out/f2_bench_1e6.json 2> gpurun_out/f2_bench_1e6.err; echo "bench1e6 rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f2_bench_ref.json 2> gpurun_out/f2_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --particles 1e6 --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 130 -c 114 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu launches rc=$?"
