#!/bin/bash
# Round 2, call K: CTA-pair kernel — operator tests, resize diagnostic, same-box A/B of the whole step with / without pairs.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_conv_halo_gpu.py tests/test_conv_train_gpu.py -m gpu -x -q > gpurun_out/pytest_conv_k.log 2>&1; echo "pytest conv rc=$?"
tail -15 gpurun_out/pytest_conv_k.log
timeout 300 python tools/resize_diag.py > gpurun_out/resize_diag.txt 2>&1; echo "resize diag rc=$?"; tail -45 gpurun_out/resize_diag.txt
timeout 900 python tools/ab_layers.py pair=0 nopair=0x10000000 > gpurun_out/ab_pair.txt 2>&1; echo "ab rc=$?"; head -5 gpurun_out/ab_pair.txt
