#!/bin/bash
# Phase A validation on one B200: moving-particle tests + the re-based bench lines (HEAD kernels).
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -k "not full_size" > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench_1e7.json 2> gpurun_out/a_bench_1e7.err
echo "bench1e7 rc=$?"
timeout 300 python bench.py --n 1e6 --steps 20 --warmup 5 --no-cpu > gpurun_out/a_bench_1e6.json 2> gpurun_out/a_bench_1e6.err
echo "bench1e6 rc=$?"
timeout 300 python bench.py --n 1e6 --steps 20 --warmup 5 --no-cpu --move 0 --no-e2e > gpurun_out/a_bench_1e6_static.json 2> gpurun_out/a_bench_1e6_static.err
echo "bench1e6 static rc=$?"
tail -c 600 gpurun_out/a_bench_1e7.json
