#!/bin/bash
# Round 2, call AE (N GPUs, N = $1): the driver's own multi-GPU command line (inference headline + data-parallel training
# legs in one run), the NCCL arithmetic check on yolov4-tiny, the training-mode bench.
set -u
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"; cut -c1-250 gpurun_out/bench_${N}gpu.json; tail -3 gpurun_out/bench_${N}gpu.err
timeout 300 $TR tools/dp_check.py > gpurun_out/dp_check_${N}gpu.jsonl 2> gpurun_out/dp_check_${N}gpu.err; echo "dp_check rc=$?"; tail -2 gpurun_out/dp_check_${N}gpu.jsonl | cut -c1-900
timeout 300 $TR tools/dp_check.py yolov4-tiny 16 > gpurun_out/dp_check_tiny_${N}gpu.jsonl 2> gpurun_out/dp_check_tiny_${N}gpu.err; echo "dp_check tiny rc=$?"; tail -2 gpurun_out/dp_check_tiny_${N}gpu.jsonl | cut -c1-700
timeout 400 $TR bench.py --gpus $N --mode train --no-cpu-baseline > gpurun_out/bench_train_${N}gpu.json 2> gpurun_out/bench_train_${N}gpu.err; echo "train ${N}gpu rc=$?"; cut -c1-250 gpurun_out/bench_train_${N}gpu.json
