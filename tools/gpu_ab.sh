# quick GPU iteration: parity tests, bench with the per-layer table, optional conv microbench
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --layers > gpurun_out/bench_a.json 2> gpurun_out/bench_layers_a.txt; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_a.json
grep -v convolutional gpurun_out/bench_layers_a.txt | tail -5
sort -k4 -n -r gpurun_out/bench_layers_a.txt | head -14
