#!/bin/bash
# Round 2, call AQ (2 GPUs): the NCCL arithmetic / replica check on the final tree (the update kernels changed since call AE).
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 240 $TR tools/dp_check.py > gpurun_out/dp_check_2gpu_final.jsonl 2> gpurun_out/dp_check_2gpu_final.err; echo "dp_check rc=$?"; tail -1 gpurun_out/dp_check_2gpu_final.jsonl | cut -c1-700
