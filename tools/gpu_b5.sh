#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_session.py tests/test_gpu_golden.py -x -q -k "not full_size" > gpurun_out/b5_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/b5_tests.log
for opts in "knn_lane=0" "knn_lane=0 tile_cfg=1" "knn_lane=0 knn_margin=0.05" "knn_lane=0 knn_margin=0.07"; do
  echo "== $opts"; python tools/knn_debug.py 1000000 0.2 $opts 2>&1 | tail -1 | cut -c1-250
done
