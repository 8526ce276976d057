#!/bin/bash
# Round 2, call AD: one-kernel tcgen05 stem as the network default — network parity, headline all-layers test, same-box
# ABAB against the operand path, bench line.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_network_gpu.py tests/test_conv_gpu.py -m gpu -x -q > gpurun_out/pytest_ad.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_ad.log
timeout 900 python -m pytest tests/test_zconfigs_gpu.py -m gpu -x -q -k "all_layers or sparse" > gpurun_out/pytest_ad2.log 2>&1; echo "pytest configs rc=$?"; tail -3 gpurun_out/pytest_ad2.log
timeout 600 python tools/ab_layers.py --batch 64 tc=0 operand=80000000 tc2=0 operand2=80000000 > gpurun_out/ab_stem_tc.txt 2>&1; echo "ab rc=$?"; head -8 gpurun_out/ab_stem_tc.txt
timeout 900 python bench.py --no-train-leg --layers > gpurun_out/bench_ad.json 2> gpurun_out/bench_layers_ad.txt; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_ad.json')); r=d['roofline']
print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(r[k],4) for k in ('achieved','frac','conv_ms_per_step','other_kernels_ms_per_step','conv_ms_per_step_bracketed','whole_step_frac')}, d['clocks'])
PY
