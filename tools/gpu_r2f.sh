#!/bin/bash
# Round 2, call F: whole suite (reproducible training, rewritten loss), bench default + train mode.
set -u
mkdir -p gpurun_out
timeout 2700 python -m pytest tests -m gpu -q -s --durations=8 > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_f.log
tail -14 gpurun_out/pytest_gpu_f.log
timeout 900 python bench.py --layers > gpurun_out/bench_f.json 2> gpurun_out/bench_layers_f.txt; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_f.json
timeout 600 python bench.py --mode train > gpurun_out/bench_train_f.json 2> gpurun_out/bench_train_f.err; echo "train rc=$?"; cut -c1-300 gpurun_out/bench_train_f.json
du -sh gpurun_out
