#!/bin/bash
mkdir -p gpurun_out
BENCH_VERBOSE=2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 4 --scaling strong --particles 5e6 --steps 6 --warmup 3 --no-e2e --no-verify > gpurun_out/s4.json 2> gpurun_out/s4.err; echo "rc=$?"
grep "^rank" gpurun_out/s4.err
