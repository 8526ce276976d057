#!/usr/bin/env python
"""Diagnostic: train step, resize there and back, train step — against the same without the detour; prints the first
tensors that differ and by how much (run on the GPU box)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N  # noqa: E402


def run(detour):
    B, hw, classes = 4, 416, 80
    text = cfgs.CONFIGS["yolov4-tiny"](batch=B, subdivisions=1, w=hw, h=hw)
    net = N.LoadSynthetic(text, cfgmod.synth_weights(text), batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    out = {}
    for s in range(2):
        if detour and s == 1:
            N.ResizeNetwork(net, 320, 352)
            N.ResizeNetwork(net, hw, hw)
        if s == 1:
            for i in range(net.n):
                if net.layer_info(i, 0) == 0:
                    for nm in ("weights", "biases", "weight_updates", "bias_updates") + (("scales", "scale_updates", "rolling_mean", "rolling_variance") if net.layer_info(i, 9) else ()):
                        try:
                            out[f"pre.L{i}.{nm}"] = N.layer_param(net, i, nm)
                        except Exception as e:  # noqa: BLE001
                            out[f"pre.L{i}.{nm}"] = np.zeros(1)
        x = cfgmod.synth_images(B, 3, hw, hw, seed=5 + s)
        truth = cfgmod.synth_truth(B, 90, classes=classes, seed=6 + s)
        out[f"cost{s}"] = np.float32(N.TrainNetworkDatum(net, x, truth))
        for i in range(net.n):
            try:
                out[f"s{s}.out.L{i}"] = net.layer_output(i)
            except Exception:  # noqa: BLE001
                pass
        for i in range(net.n):
            if net.layer_info(i, 0) == 0:
                out[f"s{s}.L{i}.weight_updates"] = N.layer_param(net, i, "weight_updates")
        N.UpdateNetwork(net)
    net.close()
    return out


def main():
    L.check(L.lib().yc_set_device(0), "yc_set_device")
    a, b = run(False), run(True)
    n = 0
    for k in a:
        if k not in b:
            print("missing", k)
            continue
        u, v = np.asarray(a[k], np.float64), np.asarray(b[k], np.float64)
        if u.shape != v.shape:
            print(f"{k}: shape {u.shape} vs {v.shape}")
            continue
        if not np.array_equal(u, v):
            d = np.abs(u - v)
            print(f"{k}: max abs diff {d.max():.3e} (max |a| {np.abs(u).max():.3e}), {int((d > 0).sum())} of {d.size} elements")
            n += 1
            if n > 40:
                break
    print("differing tensors:", n)


if __name__ == "__main__":
    main()
