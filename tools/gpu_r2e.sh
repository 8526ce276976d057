#!/bin/bash
# Round 2, call E: whole suite with the deterministic split-K wgrad and the new detection regimes, then the default bench run
# (inference headline + dense-head e2e + the training legs), then the reference arm.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
timeout 2700 python -m pytest tests -m gpu -q -s --durations=8 > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_e.log
tail -12 gpurun_out/pytest_gpu_e.log
timeout 900 python bench.py --layers > gpurun_out/bench_e.json 2> gpurun_out/bench_layers_e.txt; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_e.json
tail -5 gpurun_out/bench_layers_e.txt
timeout 600 python bench.py --mode train > gpurun_out/bench_train_e.json 2> gpurun_out/bench_train_e.err; echo "train rc=$?"; cut -c1-300 gpurun_out/bench_train_e.json
du -sh gpurun_out
