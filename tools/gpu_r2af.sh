#!/bin/bash
# Round 2, call AF: device cv::resize — bit-exact operator test, frame entry point, timing of a 64 x 1080p -> 608 batch.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_frame_resize_gpu.py tests/test_layers_gpu.py -m gpu -x -q > gpurun_out/pytest_af.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_af.log
python - <<'PY' 2>&1 | tail -5
import torch, ctypes as C, json
from darknet_b200 import lib as L
lib = L.lib()
n, sh, sw, dh, dw = 64, 1080, 1920, 608, 608
src = torch.randint(0, 256, (n, sh, sw, 3), dtype=torch.uint8, device="cuda")
dst = torch.zeros((n, 3, dh, dw), dtype=torch.float32, device="cuda")
run = lambda: L.check(lib.yc_resize_u8_to_nchw(src.data_ptr(), n, 3, sh, sw, 1, dst.data_ptr(), dh, dw, L.stream_ptr()), "resize")
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(json.dumps({"resize_1080p_to_608_x64_ms": ms, "src_mb": src.numel() / 1e6, "dst_mb": dst.numel() * 4 / 1e6,
                  "gb_per_s_src_plus_dst": (src.numel() + dst.numel() * 4) / ms / 1e6}))
PY
