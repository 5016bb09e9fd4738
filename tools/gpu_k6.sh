#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_gpu_parity.py -x -q -k "not full_size" > gpurun_out/k6_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/k6_tests.log
bash tools/gpu_k5.sh
