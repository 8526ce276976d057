#!/bin/bash
# Round 2, call M: halo patches on CTA pairs — operator tests, same-box A/B of the whole step, bench line.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_conv_halo_gpu.py -m gpu -x -q > gpurun_out/pytest_conv_m.log 2>&1; echo "pytest conv rc=$?"
tail -15 gpurun_out/pytest_conv_m.log
timeout 900 python tools/ab_layers.py pair=0 nopair=0x10000000 > gpurun_out/ab_pair_m.txt 2>&1; echo "ab rc=$?"; head -3 gpurun_out/ab_pair_m.txt; grep "128->128 @76" gpurun_out/ab_pair_m.txt | head -3
timeout 900 python bench.py --no-train-leg > gpurun_out/bench_m.json 2>/dev/null; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_m.json'))
print(d['value'], d['ms_per_step'], {k:d['roofline'][k] for k in ('achieved','frac','conv_ms_per_step','other_kernels_ms_per_step','conv_ms_per_step_bracketed','whole_step_frac')})
PY
