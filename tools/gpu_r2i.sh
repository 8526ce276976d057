#!/bin/bash
# Round 2, call I: the whole GPU suite, smoke, both bench arms, ncu launch list + full capture of 12 conv launches.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 3000 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/pytest_gpu_i.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_i.log
tail -6 gpurun_out/pytest_gpu_i.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --layers > gpurun_out/bench_i.json 2> gpurun_out/bench_layers_i.txt; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_i.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_i.json 2>/dev/null; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref_i.json
timeout 600 python bench.py --workload yolov3 --no-train-leg > gpurun_out/bench_yolov3_i.json 2>/dev/null; echo "yolov3 rc=$?"; cut -c1-300 gpurun_out/bench_yolov3_i.json
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_i.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
du -sh gpurun_out
