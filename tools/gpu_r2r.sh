#!/bin/bash
# Round 2, call R: wgrad tap groups — operator parity (both schedules), training-step tests, same-box A/B of the training bench.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_train_gpu.py tests/test_train_emul_gpu.py tests/test_repeat_gpu.py tests/test_train_gates_gpu.py -m gpu -x -q > gpurun_out/pytest_r.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_r.log
for v in groups single groups2; do
  F=""; [ $v = single ] && F="--wgrad-single-tap"
  timeout 600 python bench.py --mode train --no-cpu-baseline $F > gpurun_out/bench_train_$v.json 2>/dev/null; echo "train $v rc=$?"
  python - <<PY
import json
t=json.load(open('gpurun_out/bench_train_$v.json'))
print('$v', round(t['ms_per_step'],3), {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
done
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline > gpurun_out/bench_train_gauss.json 2>/dev/null; echo "gauss rc=$?"
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline --wgrad-single-tap > gpurun_out/bench_train_gauss_single.json 2>/dev/null
python - <<'PY'
import json
for v in ('gauss','gauss_single'):
    t=json.load(open(f'gpurun_out/bench_train_{v}.json'))
    print(v, round(t['ms_per_step'],3), {k:round(x,3) for k,x in t['roofline']['ms_by_category'].items()})
PY
