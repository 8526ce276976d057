#!/bin/bash
# Round 2, call B: operator tests first, then A/B of the specialised epilogue, then the whole suite.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py tests/test_conv_halo_gpu.py tests/test_network_gpu.py -m gpu -x -q > gpurun_out/pytest_conv.log 2>&1; echo "pytest conv rc=$?"; tail -3 gpurun_out/pytest_conv.log
timeout 900 python tools/ab_layers.py default=0 generic=0x1000000 setsv1=0x2000000 twobufs=0x800000 onebuf=0x200000 > gpurun_out/ab_layers_b.txt 2>&1; echo "ab rc=$?"
head -3 gpurun_out/ab_layers_b.txt
timeout 2400 python -m pytest tests -m gpu -q -s --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --layers --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench_layers.txt; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench.json
du -sh gpurun_out
