#!/bin/bash
# Round 2, call AJ: SPP max-pools in one launch — operator + network tests, layer table.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_layers_gpu.py tests/test_network_gpu.py -m gpu -x -q > gpurun_out/pytest_aj.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_aj.log
timeout 600 python bench.py --no-train-leg --no-cpu-baseline --layers > gpurun_out/bench_aj.json 2> gpurun_out/bench_layers_aj.txt; echo "bench rc=$?"
grep -v convolutional gpurun_out/bench_layers_aj.txt | head -12
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_aj.json')); r=d['roofline']
print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(r[k],4) for k in ('frac','conv_ms_per_step','other_kernels_ms_per_step')}, d['clocks'], d['gpu_launches'])
PY
