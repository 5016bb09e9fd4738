#!/bin/bash
mkdir -p gpurun_out
BENCH_VERBOSE=2 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --scaling strong --particles 1e7 --steps 10 --warmup 5 --no-e2e --no-verify > gpurun_out/s8.json 2> gpurun_out/s8.err; echo "rc=$?"
grep "^rank" gpurun_out/s8.err
grep "^step" gpurun_out/s8.err | tail -3
