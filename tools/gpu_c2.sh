#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --particles 1e6 --steps 10 --warmup 3 --no-strong-leg > gpurun_out/c2_bench_n2_1e6.json 2> gpurun_out/c2_bench_n2_1e6.err; echo "bench2 rc=$?"; tail -c 2500 gpurun_out/c2_bench_n2_1e6.json; tail -5 gpurun_out/c2_bench_n2_1e6.err
