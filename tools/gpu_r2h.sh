#!/bin/bash
# Round 2, call H: conv + config tests, A/B of tile pairs, bench.
set -u
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_conv_gpu.py tests/test_conv_halo_gpu.py tests/test_network_gpu.py tests/test_zconfigs_gpu.py -m gpu -q -s -k "not sparse and not dense and not plain" > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_h.log
tail -8 gpurun_out/pytest_gpu_h.log
timeout 600 python tools/ab_layers.py default=0 nopair=0x8000000 default2=0 > gpurun_out/ab_layers_h.txt 2>&1; echo "ab rc=$?"
head -3 gpurun_out/ab_layers_h.txt
du -sh gpurun_out
