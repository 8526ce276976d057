#!/bin/bash
# Round 2, call AC: tcgen05 stem — one full ncu capture (raw page + per-instruction samples exported on the box).
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stem_conv_tc -s 3 -c 1 -o /tmp/stem_tc python tools/stem_bench.py 64 608 ${1:-mish} > gpurun_out/ncu_ac.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/stem_tc.ncu-rep --page raw --csv > gpurun_out/stem_tc_raw.csv 2>/dev/null
ncu -i /tmp/stem_tc.ncu-rep --page source --csv > gpurun_out/stem_tc_source.csv 2>/dev/null
python tools/ncu_summary.py /tmp/stem_tc.ncu-rep > gpurun_out/stem_tc_summary.csv; cat gpurun_out/stem_tc_summary.csv | tail -1
