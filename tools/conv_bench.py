"""Microbenchmark of the conv kernel through the C ABI on yolov4-608 layer shapes (batch 64 by default).
Usage: python tools/conv_bench.py [--shapes all|NAME,...] [--iters N] [--batch B]
Prints one line per shape: time, TFLOP/s, GB/s of compulsory traffic, plan info."""
import argparse
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from darknet_b200 import lib as L, ops

SHAPES = {  # name: (C, N, HW, k, stride, act)
    "stem": (3, 32, 608, 3, 1, "mish"),
    "s2_32_64": (32, 64, 608, 3, 2, "mish"),
    "c1_64_64_304": (64, 64, 304, 1, 1, "mish"),
    "c3_32_64_304": (32, 64, 304, 3, 1, "mish"),
    "c1_128_64_304": (128, 64, 304, 1, 1, "mish"),
    "c3_64_64_152": (64, 64, 152, 3, 1, "mish"),
    "c3_64_32_152": (64, 32, 152, 3, 1, "mish"),
    "c3_64_96_152": (64, 96, 152, 3, 1, "mish"),
    "c3_64_128_152": (64, 128, 152, 3, 1, "mish"),
    "c1_128_128_76": (128, 128, 76, 1, 1, "mish"),
    "c3_128_128_76": (128, 128, 76, 3, 1, "mish"),
    "c3_256_256_38": (256, 256, 38, 3, 1, "mish"),
    "c1_512_256_38": (512, 256, 38, 1, 1, "leaky"),
    "c3_256_512_38": (256, 512, 38, 3, 1, "leaky"),
    "c3_512_1024_19": (512, 1024, 19, 3, 1, "leaky"),
    "c1_1024_512_19": (1024, 512, 19, 1, 1, "leaky"),
    "c3_128_256_76": (128, 256, 76, 3, 1, "leaky"),
    "head_256_255_76": (256, 255, 76, 1, 1, "linear"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="all")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--act", default=None)
    ap.add_argument("--warmup", type=int, default=3, help="untimed launches per shape (0 under ncu: one launch per shape)")
    a = ap.parse_args()
    L.check(L.lib().yc_set_device(0), "set_device")
    dev = torch.device("cuda:0")
    names = list(SHAPES) if a.shapes == "all" else a.shapes.split(",")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for nm in names:
        C, N, HW, k, s, act = SHAPES[nm]
        act = a.act or act
        B = a.batch
        ldc = (C + 7) // 8 * 8
        x = torch.randn((B, HW, HW, ldc), device=dev, dtype=torch.bfloat16)
        w = torch.randn((N, C, k, k)) * (2.0 / (C * k * k)) ** 0.5
        b = torch.zeros(N)
        conv = ops.Conv(x, (B, HW, HW, C), w, b, k, s, k // 2, act,
                        out_dtype=torch.float32 if N == 255 else torch.bfloat16)
        n, ho, wo, _ = conv.out_shape
        for _ in range(a.warmup):
            conv()
        ts = []
        for _ in range(a.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            conv()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        flops = 2.0 * B * ho * wo * N * C * k * k
        bytes_ = B * HW * HW * C * 2 + B * ho * wo * N * (4 if N == 255 else 2) + N * C * k * k * 2
        print(f"{nm:18s} {ms*1e3:9.1f} us  {flops/ms/1e9:8.1f} TF/s  {bytes_/ms/1e6:8.1f} GB/s  {conv.info()}", flush=True)
        del conv, x


if __name__ == "__main__":
    main()
