#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py tests/test_gpu_gravity.py tests/test_gpu_parity.py -x -q -k "not full_size" > gpurun_out/b6_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/b6_tests.log
python tools/knn_debug.py 1000000 0.2 2>&1 | tail -1 | cut -c1-250
python tools/knn_debug.py 1000000 0 2>&1 | tail -1 | cut -c1-250
