set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --layers > gpurun_out/bench_a.json 2> gpurun_out/bench_layers_a.txt; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_a.json
YC_MISH_RCP2=1 python bench.py --layers > gpurun_out/bench_b.json 2> gpurun_out/bench_layers_b.txt; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_b.json
python tools/conv_bench.py --shapes c1_64_64_304,c3_32_64_304,c1_128_128_76,c3_128_128_76,c3_256_256_38 > gpurun_out/cb_a.txt 2>&1
YC_MISH_RCP2=1 python tools/conv_bench.py --shapes c1_64_64_304,c3_32_64_304,c1_128_128_76,c3_128_128_76,c3_256_256_38 > gpurun_out/cb_b.txt 2>&1
cat gpurun_out/cb_a.txt gpurun_out/cb_b.txt
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s 330 -c 12 -o gpurun_out/prof_conv_early -f $CMD > gpurun_out/ncu3.log 2>&1
echo "ncu full rc=$?"
du -sh gpurun_out
