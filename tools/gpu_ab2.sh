# same-box A/B of plan variants selected by environment variables (per-layer tables under gpurun_out/)
set -u
mkdir -p gpurun_out
for i in 1 2; do
for v in 1 0 2; do
YC_W_RES=$v timeout 200 python bench.py --layers > gpurun_out/bench_w$v.json 2> gpurun_out/bench_layers_w$v.txt; python -c "import json;d=json.load(open('gpurun_out/bench_w$v.json'));print('W_RES=$v',d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])"
done
done
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
