#!/bin/bash
# Round 2, call N: ncu --set full of every conv launch of one forward (no source: the report must stay under 64 MiB),
# plus a launch list of one whole step.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain_n.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_n.log; exit 1; }
timeout 1500 ncu --set full --clock-control none -k regex:conv_fprop -c 112 -o gpurun_out/r02_conv_full -f $CMD > gpurun_out/ncu_n.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_n.csv $CMD > gpurun_out/ncu_n2.log 2>&1
echo "ncu launches rc=$?"
du -sh gpurun_out
