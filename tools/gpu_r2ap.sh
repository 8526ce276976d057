#!/bin/bash
# Round 2, call AP: ncu launch list of the headline step on the final tree (the file bench.py reads roofline.traffic from).
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed --no-train-leg"
timeout 300 $CMD > gpurun_out/plain_ap.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_ap.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_ap.csv $CMD > gpurun_out/ncu_ap.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches_ap.csv
python tools/agg_launches.py gpurun_out/launches_ap.csv > gpurun_out/step_by_kernel_ap.txt 2>&1; head -10 gpurun_out/step_by_kernel_ap.txt
