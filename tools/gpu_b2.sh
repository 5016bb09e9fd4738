#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_session.py -x -q -k "not full_size" > gpurun_out/b2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/b2_tests.log
for opts in "knn_margin=0.02" "knn_margin=0.1" "knn_margin=0.1 knn_atomic=1" "knn_margin=0.1 knn_atomic=1 cell_occ=15" "knn_margin=0.1 knn_atomic=1 cell_occ=14" "knn_margin=0.07 knn_atomic=1 cell_occ=15" "knn_margin=0.14 knn_atomic=1 cell_occ=15"; do
  echo "== $opts"; python tools/knn_debug.py 1000000 0.2 $opts 2>&1 | tail -2 | cut -c1-250
done
