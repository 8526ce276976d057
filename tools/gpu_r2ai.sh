#!/bin/bash
# Round 2, call AI: profiling evidence of the final build — ncu launch list of the headline step, one full capture of the
# stem kernel (mish), SASS opcode histogram of the library.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed --no-train-leg"
timeout 300 $CMD > gpurun_out/plain_ai.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_ai.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_ai.csv $CMD > gpurun_out/ncu_ai.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches_ai.csv
python tools/agg_launches.py gpurun_out/launches_ai.csv > gpurun_out/step_by_kernel_ai.txt 2>&1; head -12 gpurun_out/step_by_kernel_ai.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stem_conv_tc -s 3 -c 1 -o /tmp/stem_tc python tools/stem_bench.py 64 608 mish > gpurun_out/ncu_ai2.log 2>&1; echo "ncu full rc=$?"
python tools/ncu_summary.py /tmp/stem_tc.ncu-rep > gpurun_out/stem_tc_mish_summary.csv; tail -1 gpurun_out/stem_tc_mish_summary.csv | cut -c1-400
ncu -i /tmp/stem_tc.ncu-rep --page raw --csv > gpurun_out/stem_tc_mish_raw.csv 2>/dev/null
