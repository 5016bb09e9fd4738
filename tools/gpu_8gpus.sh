#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e > gpurun_out/c8b_bench_n8.json 2> gpurun_out/c8b_bench_n8.err; echo "bench8 rc=$?"; tail -c 300 gpurun_out/c8b_bench_n8.json; tail -3 gpurun_out/c8b_bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 8 --kind clustered --particles 1.25e7 --steps 5 --warmup 3 --no-e2e --no-verify --no-strong-leg > gpurun_out/c8b_bench_n8_clustered.json 2> gpurun_out/c8b_bench_n8_clustered.err; echo "clustered8 rc=$?"; tail -c 300 gpurun_out/c8b_bench_n8_clustered.json; tail -3 gpurun_out/c8b_bench_n8_clustered.err
