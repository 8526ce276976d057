#!/bin/bash
# Round 2, call P: pipelined direct stem — operator test, same-box A/B against the im2col-operand stem.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -x -q -k "stem or cta_pairs" > gpurun_out/pytest_p.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_p.log
timeout 900 python tools/ab_layers.py direct=0 im2col=0x4000 > gpurun_out/ab_stem_p.txt 2>&1; echo "ab rc=$?"; head -5 gpurun_out/ab_stem_p.txt; tail -2 gpurun_out/ab_stem_p.txt
