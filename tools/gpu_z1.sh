#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/z1_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/z1_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/z1_all.log 2>&1; echo "all rc=$?"; tail -4 gpurun_out/z1_all.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/z1_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/z1_smoke.log
