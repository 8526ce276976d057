#!/bin/bash
# One gpurun call: parity tests, bench (both arms), ncu launch list + full capture of the conv kernel.
# usage: gpurun --timeout 1800 -- 'bash tools/gpu_check.sh'
# Keep gpurun_out/ under 64 MiB (a full-set ncu capture is ~2.5 MB per launch).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
if [ "${SKIP_TESTS:-0}" = "0" ]; then
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
fi
if [ "${SKIP_BENCH:-0}" = "0" ]; then
python bench.py --layers > gpurun_out/bench.json 2> gpurun_out/bench_layers.txt; echo "bench rc=$?"; cat gpurun_out/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
fi
if [ "${SKIP_NCU:-0}" = "0" ]; then
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
# full capture (with source) of ${NCU_N:-12} conv launches starting at conv launch ${NCU_S:-372}
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s ${NCU_S:-372} -c ${NCU_N:-12} -o gpurun_out/prof_conv -f $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
fi
du -sh gpurun_out
