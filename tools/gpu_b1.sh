#!/bin/bash
# Phase B first light: parity tests of the new kernels, then a short phase timing.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/b1_parity.log 2>&1; echo "parity rc=$?" | tee -a gpurun_out/b1_parity.log
tail -15 gpurun_out/b1_parity.log
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py -x -q -k "not full_size" > gpurun_out/b1_session.log 2>&1; echo "session rc=$?" | tee -a gpurun_out/b1_session.log
tail -15 gpurun_out/b1_session.log
timeout 300 python bench.py --n 1e6 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/b1_bench_1e6.json 2> gpurun_out/b1_bench_1e6.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/b1_bench_1e6.json; tail -5 gpurun_out/b1_bench_1e6.err
