#!/bin/bash
# Round 2, call AH: the bench lines of the final build — default run (headline + training legs + layer table), the
# reference arm, configs[1] (yolov3-416), the training mode for both training configs.
set -u
mkdir -p gpurun_out
timeout 900 python bench.py --layers > gpurun_out/bench_ah.json 2> gpurun_out/bench_layers_ah.txt; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_ah.json 2> gpurun_out/bench_ref_ah.err; echo "reference rc=$?"; cut -c1-300 gpurun_out/bench_ref_ah.json
timeout 600 python bench.py --workload yolov3 --no-train-leg --no-cpu-baseline > gpurun_out/bench_yolov3_ah.json 2> gpurun_out/bench_yolov3_ah.err; echo "yolov3 rc=$?"
timeout 600 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_ah.json 2> gpurun_out/bench_train_ah.err; echo "train rc=$?"
timeout 600 python bench.py --mode train --train-net gaussian-yolov3 --batch 8 --no-cpu-baseline > gpurun_out/bench_train_g_ah.json 2> gpurun_out/bench_train_g_ah.err; echo "train gaussian rc=$?"
python - <<'PY'
import json
for f in ('bench_ah','bench_yolov3_ah','bench_train_ah','bench_train_g_ah'):
    try:
        d=json.load(open(f'gpurun_out/{f}.json')); r=d['roofline']
        print(f, round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), {k:(round(r[k],4) if isinstance(r.get(k),(int,float)) else r.get(k)) for k in ('achieved','frac','conv_ms_per_step','other_kernels_ms_per_step','traffic')}, d['clocks'], d.get('gpu_launches'))
    except Exception as e: print(f, 'ERR', e)
PY
