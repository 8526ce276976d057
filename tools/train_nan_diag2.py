#!/usr/bin/env python
"""Diagnostic: bench.py --mode train's own sequence (device-resident steps enqueued back to back on torch's stream) with
host synchronisation at different points: every variant must end at the same cost (the step is deterministic)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N  # noqa: E402

L.check(L.lib().yc_set_device(0), "yc_set_device")
torch.cuda.set_device(0)
name, hw, B = "yolov4-tiny", 416, 64
text = cfgs.CONFIGS[name](w=hw, h=hw, batch=B, subdivisions=1)
weights = cfgmod.synth_weights(text)
x = cfgmod.synth_images(B, 3, hw, hw, seed=1234)
truth = cfgmod.synth_truth(B, 90, classes=80, seed=4321)
xd, td = torch.from_numpy(x).cuda(), torch.from_numpy(truth).cuda()


def run(mode, steps=25, side=False):
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    s2 = torch.cuda.Stream() if side else None
    ctx = torch.cuda.stream(s2) if side else torch.cuda.stream(torch.cuda.current_stream())
    with ctx:
        sp = L.stream_ptr()
        for s in range(steps):
            N.ForwardBackwardNetworkGpu(net, xd.data_ptr(), td.data_ptr(), sp)
            if mode in ("fb", "both"):
                torch.cuda.synchronize()
            N.UpdateNetwork(net, stream=sp)
            if mode in ("upd", "both"):
                torch.cuda.synchronize()
        torch.cuda.synchronize()
    c = float(N.TrainNetworkDatum(net, x, truth))
    w = np.concatenate([N.layer_param(net, i, "weights") for i in range(net.n) if net.layer_info(i, 0) == 0])
    net.close()
    return c, float(np.abs(w).sum()), bool(np.isfinite(w).all())


for mode in ("both", "fb", "upd", "none", "none", "both"):
    print(mode, run(mode), flush=True)
print("side stream none", run("none", side=True), flush=True)
