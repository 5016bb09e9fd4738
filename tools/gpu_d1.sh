#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/d1_bench_1e7.json 2> gpurun_out/d1_bench_1e7.err; echo "bench1e7 rc=$?"; tail -3 gpurun_out/d1_bench_1e7.err
timeout 600 python bench.py --particles 4e6 --kind clustered --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/d1_bench_cl4e6.json 2> gpurun_out/d1_bench_cl4e6.err; echo "clustered rc=$?"; tail -3 gpurun_out/d1_bench_cl4e6.err
timeout 1200 python -m pytest tests -m gpu -x -q -k "not full_size_steps" > gpurun_out/d1_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/d1_pytest.log
