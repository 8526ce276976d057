bash tools/gpu_check.sh
python bench.py --mode train > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "train rc=$?"; cut -c1-300 gpurun_out/bench_train.json
python bench.py --mode train --impl reference --steps 2 --warmup 1 > gpurun_out/bench_train_ref.json 2>/dev/null; echo "train ref rc=$?"; cut -c1-200 gpurun_out/bench_train_ref.json
