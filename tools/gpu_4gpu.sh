set -u
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_4gpu.json 2> gpurun_out/bench_4gpu.err; echo "rc=$?"; cut -c1-250 gpurun_out/bench_4gpu.json
