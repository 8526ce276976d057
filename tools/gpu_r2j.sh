#!/bin/bash
# Round 2, call J: the no-alloc entry point's test alone in a fresh process, then the whole GPU suite (the same test
# runs again after all others), smoke, the bench line with the per-layer table.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 3000 python -m pytest tests -m gpu -x -q --durations=25 > gpurun_out/pytest_gpu_j.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_j.log
tail -40 gpurun_out/pytest_gpu_j.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --layers > gpurun_out/bench_j.json 2> gpurun_out/bench_layers_j.txt; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_j.json
