# two-GPU bench lines (inference: image-sharded, no collective; training: data-parallel NCCL all-reduce)
set -u
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "rc=$?"; cut -c1-250 gpurun_out/bench_2gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --mode train --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_train_2gpu.json 2> gpurun_out/bench_train_2gpu.err; echo "rc=$?"; cut -c1-250 gpurun_out/bench_train_2gpu.json
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1gpu_samebox.json 2>/dev/null; cut -c1-200 gpurun_out/bench_1gpu_samebox.json
