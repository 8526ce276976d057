#!/usr/bin/env python
"""Diagnostic: N training steps on a fixed batch, cost per step; two fresh instances per setting (bit reproducibility over
many steps), both wgrad schedules; reports the first layer whose parameters / gradients stop being finite."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N  # noqa: E402


def run(name, hw, B, steps, tapgroups, check_layers):
    L.lib().yc_conv_wgrad_set_tap_groups(tapgroups)
    text = cfgs.CONFIGS[name](w=hw, h=hw, batch=B, subdivisions=1)
    net = N.LoadSynthetic(text, cfgmod.synth_weights(text), batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    x = cfgmod.synth_images(B, 3, hw, hw, seed=1234)
    truth = cfgmod.synth_truth(B, 90, classes=80, seed=4321)
    convs = [i for i in range(net.n) if net.layer_info(i, 0) == 0]
    costs = []
    bad = None
    for s in range(steps):
        c = N.TrainNetworkDatum(net, x, truth)
        costs.append(float(c))
        if check_layers and bad is None:
            for i in convs:
                for key in ("weight_updates", "weights") + (("mean", "variance", "scales", "biases") if net.layer_info(i, 9) else ("biases",)):
                    g = N.layer_param(net, i, key)
                    if not np.isfinite(g).all():
                        bad = (s, i, key, int((~np.isfinite(g)).sum()), net.layer_info(i, 5), net.layer_info(i, 1), net.layer_info(i, 6))
                        break
                    if key == "variance" and (g < 1e-12).any():
                        print(f"   step {s} layer {i}: {int((g < 1e-12).sum())} channels with zero batch variance, max |w| {np.abs(N.layer_param(net, i, 'weights')).max():.3e}")
                if bad:
                    break
        N.UpdateNetwork(net)
        if not np.isfinite(c):
            break
    lr = L.lib().yc_network_current_rate(net.h, -1, 0) if hasattr(L.lib(), "yc_network_current_rate") else None
    net.close()
    return costs, bad, lr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--nets", default="yolov4-tiny,gaussian-yolov3")
    a = ap.parse_args()
    L.check(L.lib().yc_set_device(0), "yc_set_device")
    for name, hw, B in (("yolov4-tiny", 416, 64), ("gaussian-yolov3", 608, 8)):
        if name not in a.nets.split(","):
            continue
        for tg in (1, 0):
            r1 = run(name, hw, B, a.steps, tg, True)
            r2 = run(name, hw, B, a.steps, tg, False)
            same = r1[0] == r2[0] or (len(r1[0]) == len(r2[0]) and all((u == v) or (u != u and v != v) for u, v in zip(r1[0], r2[0])))
            print(f"{name} tapgroups={tg}: costs {['%.4f' % c for c in r1[0][:6]]} ... {['%.4f' % c for c in r1[0][-3:]]}  "
                  f"steps={len(r1[0])} identical_across_instances={same} first_nonfinite={r1[1]} lr={r1[2]}", flush=True)
            if not same:
                k = next(i for i, (u, v) in enumerate(zip(r1[0], r2[0])) if u != v and not (u != u and v != v))
                print(f"   first differing step {k}: {r1[0][k]!r} vs {r2[0][k]!r}")


if __name__ == "__main__":
    main()
