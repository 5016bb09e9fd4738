#!/bin/bash
# usage: gpu_c4.sh NGPUS
N=$1
mkdir -p gpurun_out
if [ "$N" = "4" ]; then
  timeout 600 python -m pytest tests/test_gpu_domain.py -x -q > gpurun_out/c4_domain.log 2>&1; echo "domain rc=$?"; tail -4 gpurun_out/c4_domain.log
fi
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/c4_bench_n$N.json 2> gpurun_out/c4_bench_n$N.err; echo "bench$N rc=$?"; tail -c 600 gpurun_out/c4_bench_n$N.json; tail -4 gpurun_out/c4_bench_n$N.err
