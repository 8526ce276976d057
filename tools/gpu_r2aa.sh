#!/bin/bash
# Round 2, call AA: widened CTA-pair rule — conv operator tests, whole-net parity at the headline shapes, bench line.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_conv_halo_gpu.py tests/test_network_gpu.py -m gpu -x -q > gpurun_out/pytest_aa.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_aa.log
timeout 900 python -m pytest tests/test_zconfigs_gpu.py -m gpu -x -q -k "all_layers or sparse" > gpurun_out/pytest_aa2.log 2>&1; echo "pytest configs rc=$?"; tail -3 gpurun_out/pytest_aa2.log
timeout 900 python bench.py --no-train-leg --layers > gpurun_out/bench_aa.json 2> gpurun_out/bench_layers_aa.txt; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_aa.json')); r=d['roofline']
print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(r[k],4) for k in ('achieved','frac','conv_ms_per_step','other_kernels_ms_per_step','conv_ms_per_step_bracketed','whole_step_frac')}, d['clocks'])
PY
