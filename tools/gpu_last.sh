set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py --layers > gpurun_out/bench.json 2> gpurun_out/bench_layers.txt; echo "bench rc=$?"; cut -c1-160 gpurun_out/bench.json
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s 330 -c 12 -o gpurun_out/prof_conv_early_final -f $CMD > gpurun_out/ncu4.log 2>&1
echo "ncu rc=$?"
