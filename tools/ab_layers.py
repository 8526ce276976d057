#!/usr/bin/env python
"""Same-box A/B of plan switches: load the workload once per flag set, time the graph-replayed step and every layer
(yc_network_profile), print the tables side by side.  Boxes of the pool differ by +-4 % under the power cap, so kernel
variants are only ever compared inside one run of this tool.

usage: python tools/ab_layers.py [--workload yolov4] [--batch 64] name=hexflags ...
       (flags are OR'ed onto YC_NET_DEFAULT & ~PULL_HEADS; plan switches go in bits 16..23, e.g. onebuf=0x200000)"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="yolov4")
    ap.add_argument("--hw", type=int, default=608)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("sets", nargs="*", default=["default=0"])
    args = ap.parse_args()
    L.check(L.lib().yc_set_device(0), "yc_set_device")
    torch.cuda.set_device(0)
    text = cfgs.CONFIGS[args.workload](w=args.hw, h=args.hw)
    weights = cfgmod.synth_weights(text)
    cfgmod.shift_head_objectness(text, weights, -6.0, 40.0)
    x = torch.from_numpy(cfgmod.synth_images(args.batch, 3, args.hw, args.hw)).cuda()
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    cols, names, steps, meta = [], [], [], None
    for spec in args.sets:
        name, _, hx = spec.partition("=")
        extra = int(hx or "0", 16)
        L.lib().yc_conv_fprop_set_issue_v2(0 if name.startswith("issue1") else 1)   # (process-wide A/B switch)
        net = N.LoadSynthetic(text, weights, batch=args.batch, flags=(L.NET_DEFAULT & ~L.NET_PULL_HEADS) | extra)
        for _ in range(3):
            L.check(L.lib().yc_network_forward(net.h, C.c_void_p(x.data_ptr()), sp), "forward")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            L.check(L.lib().yc_network_forward(net.h, C.c_void_p(x.data_ptr()), sp), "forward")
        e1.record()
        torch.cuda.synchronize()
        steps.append(e0.elapsed_time(e1) / args.iters)
        ms = np.zeros(net.n + 1, np.float32)
        L.check(L.lib().yc_network_profile(net.h, C.c_void_p(x.data_ptr()), args.iters, ms.ctypes.data_as(C.c_void_p)), "profile")
        info = [(net.layer_type(i), [net.layer_info(i, k) for k in (6, 7, 5, 1, 2, 10, 11, 12, 13, 14)]) for i in range(net.n)]
        cols.append((ms, info))
        names.append(name)
        net.close()
        del net
        torch.cuda.empty_cache()
    print("step ms (graph replay): " + "  ".join(f"{n}={s:.3f}" for n, s in zip(names, steps)))
    conv = [sum(float(ms[i]) for i in range(len(info)) if info[i][0] == "convolutional") for ms, info in cols]
    print("conv ms (event-bracketed): " + "  ".join(f"{n}={s:.3f}" for n, s in zip(names, conv)))
    n = len(cols[0][1])
    print(f"{'layer':>5} {'geometry':28s} " + " ".join(f"{nm:>22s}" for nm in names))
    for i in range(n):
        if not any(c[0][i] > 0 for c in cols):
            continue
        t, (k, s, ci, co, ho, *_rest) = cols[0][1][i]
        geo = f"k{k} s{s} {ci}->{co} @{ho}" if t == "convolutional" else t
        cells = []
        for ms, info in cols:
            bn, st, tiles, var, pf = info[i][1][5:10]
            cells.append(f"{ms[i]*1e3:8.1f} bn{bn:<3d} v{var:<3d}" + (f" <L{pf}" if pf >= 0 else "     "))
        print(f"L{i:<4d} {geo:28s} " + " ".join(f"{c:>22s}" for c in cells))
    print("input convert " + " ".join(f"{c[0][n]*1e3:8.1f}" for c in cols))


if __name__ == "__main__":
    main()
