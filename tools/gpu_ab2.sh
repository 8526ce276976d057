set -u
mkdir -p gpurun_out
cp darknet_b200/libyc_b200.so /tmp/new.so
for i in 1 2; do
cp /tmp/new.so darknet_b200/libyc_b200.so
timeout 200 python bench.py > gpurun_out/bench_new$i.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_new$i.json'));print('new',d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])"
cp darknet_b200/libyc_b200_prev.so darknet_b200/libyc_b200.so
timeout 200 python bench.py > gpurun_out/bench_prev$i.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_prev$i.json'));print('prev',d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])"
done
cp /tmp/new.so darknet_b200/libyc_b200.so
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 200 python bench.py --mode train > gpurun_out/bench_train_new.json 2>/dev/null; cut -c1-160 gpurun_out/bench_train_new.json
