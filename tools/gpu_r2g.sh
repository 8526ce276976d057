#!/bin/bash
# Round 2, call G: the two failing config tests + conv tests, A/B of the Newton-reciprocal mish, bench.
set -u
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_conv_gpu.py tests/test_zconfigs_gpu.py -m gpu -q -s -k "conv or yolov3 or gaussian or default_plan" > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_g.log
tail -8 gpurun_out/pytest_gpu_g.log
timeout 600 python tools/ab_layers.py default=0 mufu=0x4000000 > gpurun_out/ab_layers_g.txt 2>&1; echo "ab rc=$?"
head -3 gpurun_out/ab_layers_g.txt
timeout 600 python bench.py --layers --no-train-leg > gpurun_out/bench_g.json 2> gpurun_out/bench_layers_g.txt; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_g.json
du -sh gpurun_out
