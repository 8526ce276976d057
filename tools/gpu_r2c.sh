#!/bin/bash
# Round 2, call C: whole -m gpu suite (shim, a17, reworked detection identity), A/B, ncu full capture of the first
# twenty conv launches of a step with the specialised epilogue.
set -u
mkdir -p gpurun_out
timeout 2700 python -m pytest tests -m gpu -q -s --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python tools/ab_layers.py default=0 twobufs=0x800000 generic=0x1000000 > gpurun_out/ab_layers_c.txt 2>&1; echo "ab rc=$?"
head -3 gpurun_out/ab_layers_c.txt
timeout 300 python bench.py --layers --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench_layers.txt; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench.json
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s 315 -c 20 -o gpurun_out/prof_conv_c -f $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
du -sh gpurun_out
