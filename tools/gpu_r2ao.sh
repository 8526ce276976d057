#!/bin/bash
# Round 2, call AO: where the stride-2 32->64 layer (pixel-pair view, im2col path) waits — one full ncu capture of the
# first conv_fprop_kernel_t launch of a forward pass (= layer 1), with per-instruction stall samples.
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_fprop_kernel_t -s 0 -c 1 -o /tmp/l1 python bench.py --steps 1 --warmup 1 --quick-exit-after-timed --no-train-leg > gpurun_out/ncu_ao.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/l1.ncu-rep --page raw --csv > gpurun_out/l1_raw.csv 2>/dev/null
ncu -i /tmp/l1.ncu-rep --page source --csv | awk -F'","' 'NR<3 || $6+0 > 0' > gpurun_out/l1_source.csv 2>/dev/null
ls -la gpurun_out/l1_*.csv; python tools/ncu_summary.py /tmp/l1.ncu-rep | tail -1 | cut -c1-300
