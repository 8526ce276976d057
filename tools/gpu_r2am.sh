#!/bin/bash
# Round 2, call AM: where the 3x3 32->64 @304 halo layer waits — one full ncu capture with per-instruction stall samples.
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_fprop -c 1 -o /tmp/c3 python tools/conv_bench.py --shapes c3_32_64_304 --warmup 0 --iters 1 > gpurun_out/ncu_am.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/c3.ncu-rep --page raw --csv > gpurun_out/c3_raw.csv 2>/dev/null
ncu -i /tmp/c3.ncu-rep --page source --csv > gpurun_out/c3_source.csv 2>/dev/null
ls -la gpurun_out/c3_*.csv; tail -3 gpurun_out/ncu_am.log
