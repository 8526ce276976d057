#!/bin/bash
# Round 2, call U: batchnorm finalize / backward coefficients folded into their producers — training tests, training bench.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_conv_train_gpu.py tests/test_train_emul_gpu.py tests/test_train_gpu.py tests/test_repeat_gpu.py tests/test_train_gates_gpu.py tests/test_resize_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_u.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_u.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
for v in a b; do
  timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_u_$v.json 2>/dev/null; echo "train rc=$?"
  python - <<PY
import json
t=json.load(open('gpurun_out/bench_train_u_$v.json'))
print(round(t['ms_per_step'],3), t['roofline']['frac'], {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
done
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline > gpurun_out/bench_train_u_gauss.json 2>/dev/null
python - <<'PY'
import json
t=json.load(open('gpurun_out/bench_train_u_gauss.json'))
print('gauss', round(t['ms_per_step'],3), t['roofline']['frac'], {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
