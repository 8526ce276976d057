#!/bin/bash
# Round 2, call T (2 GPUs): the driver's own multi-GPU command line (inference headline + data-parallel training legs in one
# run), the NCCL arithmetic check, the training-mode bench.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench 2gpu rc=$?"; cut -c1-250 gpurun_out/bench_2gpu.json; tail -3 gpurun_out/bench_2gpu.err
timeout 600 $TR tools/dp_check.py > gpurun_out/dp_check_2gpu.jsonl 2> gpurun_out/dp_check_2gpu.err; echo "dp_check rc=$?"; tail -5 gpurun_out/dp_check_2gpu.jsonl
timeout 600 $TR bench.py --gpus 2 --mode train --no-cpu-baseline > gpurun_out/bench_train_2gpu.json 2> gpurun_out/bench_train_2gpu.err; echo "train 2gpu rc=$?"; cut -c1-250 gpurun_out/bench_train_2gpu.json
