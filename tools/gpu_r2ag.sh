#!/bin/bash
# Round 2, call AG: the driver's round-end sequence on the final build — full GPU suite, smoke(), and which .so files the
# python processes loaded.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_ag.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_ag.log
timeout 300 python -c "
import __graft_entry__ as g
g.smoke()
import os
print('loaded:', sorted({l.split()[-1] for l in open('/proc/self/maps') if 'libyc' in l or 'yolo_core_b200' in l}))
" > gpurun_out/smoke_ag.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke_ag.log
