#!/bin/bash
# Round 2, call Z: YOLO decode fused into the head convolutions — operator test, whole-net parity incl. detections, A/B.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_layers_gpu.py tests/test_network_gpu.py tests/test_golden.py tests/test_shim_gpu.py -m gpu -x -q > gpurun_out/pytest_z.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_z.log
timeout 1200 python -m pytest tests/test_zconfigs_gpu.py -m gpu -x -q -k "sparse or yolov3_416_b64 or gaussian_yolov3_608_detections" > gpurun_out/pytest_z2.log 2>&1; echo "pytest configs rc=$?"; tail -4 gpurun_out/pytest_z2.log
timeout 900 python tools/ab_layers.py --iters 10 fused=0 separate=0x200000 fused2=0 separate2=0x200000 > gpurun_out/ab_decode.txt 2>&1; echo "ab rc=$?"; head -3 gpurun_out/ab_decode.txt; grep -E "^L(138|139|149|150|160|161) " gpurun_out/ab_decode.txt
