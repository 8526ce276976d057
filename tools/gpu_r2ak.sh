#!/bin/bash
# Round 2, call AK: the driver's round-end sequence on the final tree — full GPU suite, smoke(), bench line + reference arm.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_ak.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_ak.log
timeout 300 python -c "
import __graft_entry__ as g
g.smoke()
print('loaded:', sorted({l.split()[-1] for l in open('/proc/self/maps') if 'libyc' in l or 'yolo_core_b200' in l}))
" > gpurun_out/smoke_ak.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke_ak.log
timeout 900 python bench.py --layers > gpurun_out/bench_ak.json 2> gpurun_out/bench_layers_ak.txt; echo "bench rc=$?"
timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_ak.json 2> gpurun_out/bench_train_ak.err; echo "train rc=$?"
timeout 600 python bench.py --workload yolov3 --no-train-leg --no-cpu-baseline > gpurun_out/bench_yolov3_ak.json 2> /dev/null; echo "yolov3 rc=$?"
python - <<'PY'
import json
for f in ('bench_ak','bench_train_ak','bench_yolov3_ak'):
    d=json.load(open(f'gpurun_out/{f}.json')); r=d['roofline']
    print(f, round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), {k:(round(r[k],4) if isinstance(r.get(k),(int,float)) else r.get(k)) for k in ('achieved','frac','conv_ms_per_step','other_kernels_ms_per_step','traffic')}, d['clocks'], d.get('gpu_launches'))
PY
