#!/bin/bash
# Round 2, call W: NULL-stream fix of the training entry points — regression test, training benches on an explicit stream.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_repeat_gpu.py tests/test_train_gates_gpu.py -m gpu -x -q > gpurun_out/pytest_w.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_w.log
timeout 600 python tools/train_nan_diag2.py > gpurun_out/train_nan_diag5.txt 2>&1; tail -8 gpurun_out/train_nan_diag5.txt
timeout 600 python bench.py --mode train > gpurun_out/bench_train_w.json 2>/dev/null; echo "train rc=$?"
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline > gpurun_out/bench_train_w_gauss.json 2>/dev/null; echo "gauss rc=$?"
python - <<'PY'
import json
for v in ('w','w_gauss'):
    t=json.load(open(f'gpurun_out/bench_train_{v}.json'))
    print(v, round(t['ms_per_step'],3), round(t['value'],1), 'e2e', round(t['e2e']['value'],1), 'cost', t['e2e']['last_cost'], round(t['roofline']['frac'],4), {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
