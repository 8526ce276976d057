#!/bin/bash
# Round 2, call Q: whole GPU suite, smoke, bench (headline + per-layer table), training-mode bench with its breakdown,
# yolov3-416 line, reference arm.
set -u
mkdir -p gpurun_out
timeout 3000 python -m pytest tests -m gpu -q --durations=10 > gpurun_out/pytest_gpu_q.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_q.log
tail -18 gpurun_out/pytest_gpu_q.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --layers > gpurun_out/bench_q.json 2> gpurun_out/bench_layers_q.txt; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_q.json
timeout 900 python bench.py --mode train > gpurun_out/bench_train_q.json 2> gpurun_out/bench_train_q.err; echo "bench train rc=$?"; cut -c1-200 gpurun_out/bench_train_q.json
timeout 600 python bench.py --workload yolov3 --no-train-leg > gpurun_out/bench_yolov3_q.json 2>/dev/null; echo "yolov3 rc=$?"; cut -c1-200 gpurun_out/bench_yolov3_q.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_q.json 2>/dev/null; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_ref_q.json
