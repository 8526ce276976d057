#!/bin/bash
# Round 2, call X (re-used): training tests + training benches — training tests, training benches.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_conv_train_gpu.py tests/test_train_emul_gpu.py tests/test_train_gpu.py tests/test_repeat_gpu.py tests/test_train_gates_gpu.py tests/test_resize_gpu.py tests/test_zconfigs_gpu.py -m gpu -x -q -k "not detections" > gpurun_out/pytest_x.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_x.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_x.json 2>/dev/null; echo "train rc=$?"
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline > gpurun_out/bench_train_x_gauss.json 2>/dev/null; echo "gauss rc=$?"
python - <<'PY'
import json
for v in ('x','x_gauss'):
    t=json.load(open(f'gpurun_out/bench_train_{v}.json'))
    print(v, round(t['ms_per_step'],3), round(t['value'],1), 'e2e', round(t['e2e']['value'],1), 'cost', t['e2e']['last_cost'], round(t['roofline']['frac'],4), {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
