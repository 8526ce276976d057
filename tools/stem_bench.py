"""Time the stem convolution entry points alone (3 -> 32, 3x3, mish) on a yolov4-608 x 64 input:
   python tools/stem_bench.py [batch] [hw]
CUDA events on the launching stream, inputs (94 MB fp32) + outputs (1.5 GB bf16) far larger than L2."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from darknet_b200 import lib as L


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    HW = int(sys.argv[2]) if len(sys.argv) > 2 else 608
    lib = L.lib()
    gpu = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(1)
    x = torch.rand((B, 3, HW, HW), generator=g).to(gpu)
    w2 = (torch.randn((32, 32), generator=g) * 0.27).to(gpu)
    w2[:, 27:] = 0
    packed = torch.zeros(32 * 32, dtype=torch.bfloat16, device=gpu)
    L.check(lib.yc_conv_pack(w2.data_ptr(), 32, 32, 1, packed.data_ptr(), L.stream_ptr()), "pack")
    bias = (torch.rand(32, generator=g) - 0.5).to(gpu)
    out = {}
    ys = {}
    act = sys.argv[3] if len(sys.argv) > 3 else "mish"
    out["act"] = act
    for entry in ("yc_stem_conv_direct", "yc_stem_conv_tc"):
        y = torch.zeros((B, HW, HW, 32), dtype=torch.bfloat16, device=gpu)
        view = L.view(y, B, HW, HW, 32, 32)
        fn = getattr(lib, entry)
        run = lambda: L.check(fn(x.data_ptr(), B, HW, HW, packed.data_ptr(), bias.data_ptr(), L.ACT[act], view, L.stream_ptr()), entry)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            run()
        e1.record()
        torch.cuda.synchronize()
        out[entry + "_ms"] = e0.elapsed_time(e1) / 10
        ys[entry] = y
    a, b = ys["yc_stem_conv_direct"].float(), ys["yc_stem_conv_tc"].float()
    out["max_abs_diff_between_entries"] = float((a - b).abs().max())
    out["gb_per_s_tc"] = (B * HW * HW * (12 + 64)) / out["yc_stem_conv_tc_ms"] / 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
