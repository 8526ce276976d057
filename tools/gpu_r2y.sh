#!/bin/bash
# Round 2, call Y: ncu launch list of one training step (yolov4-tiny 416 b64) -> per-kernel and per-launch times.
set -u
mkdir -p gpurun_out
CMD="python bench.py --mode train --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain_y.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_y.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_train_y.csv $CMD > gpurun_out/ncu_y.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches_train_y.csv
