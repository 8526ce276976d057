#!/bin/bash
# Round 2, call D: the config-level tests again (logs of call C were lost: gpurun_out over 64 MiB), the training tests with the
# new default semantics, ncu full capture of twelve conv launches (stem .. 3x3 64->64 @152).
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_zconfigs_gpu.py tests/test_train_emul_gpu.py tests/test_train_gpu.py tests/test_shim_gpu.py tests/test_layers_gpu.py -m gpu -q -s --durations=10 > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_d.log
tail -12 gpurun_out/pytest_gpu_d.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
CMD="python bench.py --steps 1 --warmup 3 --quick-exit-after-timed"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:conv_fprop -s 315 -c 12 -o gpurun_out/prof_conv_d -f $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
du -sh gpurun_out
