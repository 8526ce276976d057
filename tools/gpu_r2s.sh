#!/bin/bash
# Round 2, call S: programmatic dependent launch of the conv kernels — whole-net parity, same-box A/B.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_network_gpu.py tests/test_u8_input_gpu.py tests/test_shim_gpu.py -m gpu -x -q > gpurun_out/pytest_s.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_s.log
timeout 900 python -m pytest tests/test_zconfigs_gpu.py -m gpu -x -q -k "all_layers or sparse" > gpurun_out/pytest_s2.log 2>&1; echo "pytest configs rc=$?"; tail -4 gpurun_out/pytest_s2.log
timeout 900 python tools/ab_layers.py --iters 20 pdl=0 nopdl=0x8000 pdl2=0 nopdl2=0x8000 > gpurun_out/ab_pdl.txt 2>&1; echo "ab rc=$?"; head -3 gpurun_out/ab_pdl.txt
