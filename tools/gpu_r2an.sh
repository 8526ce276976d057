#!/bin/bash
# Round 2, call AN: unrolled halo tap issue (resident weights) — halo operator tests, same-box ABAB per layer.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_halo_gpu.py tests/test_conv_gpu.py tests/test_conv_train_gpu.py -m gpu -x -q > gpurun_out/pytest_an.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_an.log
timeout 600 python tools/ab_layers.py --batch 64 v2=0 issue1=0 v2b=0 issue1b=0 > gpurun_out/ab_issue.txt 2>&1; echo "ab rc=$?"; head -4 gpurun_out/ab_issue.txt | cut -c1-200; grep -E "^L(1|2|5|8|10|11|24|54|55|85|86) " gpurun_out/ab_issue.txt | cut -c1-200
