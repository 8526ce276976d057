#!/bin/bash
# Round 2, call L: whole GPU suite (no -x: every failure shows), smoke, bench with the per-layer table.
set -u
mkdir -p gpurun_out
timeout 3000 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/pytest_gpu_l.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_l.log
tail -30 gpurun_out/pytest_gpu_l.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --layers > gpurun_out/bench_l.json 2> gpurun_out/bench_layers_l.txt; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_l.json
