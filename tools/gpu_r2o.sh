#!/bin/bash
# Round 2, call O: direct stem + halo CTA pairs (operator tests, whole-net parity at the headline shapes, same-box A/B),
# ncu --set full of ONE launch per shape class (exported to CSV on the box: the report itself may not fit the 64 MiB
# return channel), launch list of one whole step.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_network_gpu.py -m gpu -x -q > gpurun_out/pytest_o.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_o.log
timeout 900 python -m pytest tests/test_zconfigs_gpu.py -m gpu -x -q -k "all_layers or sparse" > gpurun_out/pytest_o2.log 2>&1; echo "pytest configs rc=$?"
tail -4 gpurun_out/pytest_o2.log
timeout 900 python tools/ab_layers.py direct=0 im2col=0x4000 > gpurun_out/ab_stem.txt 2>&1; echo "ab rc=$?"; head -5 gpurun_out/ab_stem.txt; tail -2 gpurun_out/ab_stem.txt
SH=c3_512_1024_19,c3_256_256_38,c3_128_128_76,c1_1024_512_19,c1_512_256_38,c1_128_128_76,c1_64_64_304,c3_32_64_304,c3_64_64_152,s2_32_64
timeout 600 python tools/conv_bench.py --shapes $SH --iters 1 --warmup 0 > gpurun_out/conv_bench_o.txt 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:conv_fprop -c 10 -o gpurun_out/r02_conv_classes -f python tools/conv_bench.py --shapes $SH --iters 1 --warmup 0 > gpurun_out/ncu_o.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
python tools/ncu_summary.py gpurun_out/r02_conv_classes.ncu-rep > gpurun_out/r02_conv_classes_ncu_full.csv 2> gpurun_out/ncu_summary.err
SZ=$(stat -c %s gpurun_out/r02_conv_classes.ncu-rep 2>/dev/null || echo 0)
if [ "$SZ" -gt 45000000 ]; then echo "report $SZ bytes: too large to return, keeping the CSV only"; rm -f gpurun_out/r02_conv_classes.ncu-rep; fi
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_o.csv $CMD > gpurun_out/ncu_o2.log 2>&1
echo "ncu launches rc=$?"
du -sh gpurun_out
