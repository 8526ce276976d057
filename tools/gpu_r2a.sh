#!/bin/bash
# Round 2, call A: full -m gpu suite (new config-parity tests included), same-box A/B of the new plan switches,
# one ncu --set full capture of the first eight conv launches of a step.
# usage: gpurun --timeout 2400 -- 'bash tools/gpu_r2a.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q -s --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/ab_layers.py default=0 onebuf=0x200000 costv1=0x400000 nopair=0x1000 > gpurun_out/ab_layers.txt 2>&1; echo "ab rc=$?"
head -3 gpurun_out/ab_layers.txt
timeout 300 python bench.py --layers --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench_layers.txt; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s ${NCU_S:-315} -c ${NCU_N:-8} -o gpurun_out/prof_conv -f $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
du -sh gpurun_out
