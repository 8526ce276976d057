#!/bin/bash
# Round 2, call AB: tcgen05 stem kernel — operator parity against the oracle, then timing against the mma.sync kernel.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -x -q -k "stem_conv_direct" > gpurun_out/pytest_ab.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_ab.log
timeout 300 python tools/stem_bench.py > gpurun_out/stem_bench_ab.json 2> gpurun_out/stem_bench_ab.err; echo "stem bench rc=$?"; cat gpurun_out/stem_bench_ab.json; tail -5 gpurun_out/stem_bench_ab.err
