#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c1_gpus.txt
timeout 900 python -m pytest tests/test_gpu_domain.py -x -q > gpurun_out/c1_domain.log 2>&1; echo "domain rc=$?"; tail -25 gpurun_out/c1_domain.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --particles 1e6 --steps 10 --warmup 3 > gpurun_out/c1_bench_n2_1e6.json 2> gpurun_out/c1_bench_n2_1e6.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/c1_bench_n2_1e6.json; tail -5 gpurun_out/c1_bench_n2_1e6.err
