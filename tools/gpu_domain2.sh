#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_domain.py -x -q -m gpu > gpurun_out/d2_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/d2_tests.log
