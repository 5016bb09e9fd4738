#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/numa_probe.py > gpurun_out/g8_numa.log 2>&1; echo "probe rc=$?"; tail -40 gpurun_out/g8_numa.log
