"""Run-to-run determinism of the training step: two fresh networks in one process (the second reuses the first's freed
device memory) must give the same gradients.  python tools/repeat_check.py [micro|yolov4-tiny] [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "micro"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    hw = 64 if name == "micro" else 416
    classes = 3 if name == "micro" else 80
    kw = {} if name == "micro" else dict(w=hw, h=hw)
    text = cfgs.CONFIGS[name](batch=B, subdivisions=1, **kw)
    weights = cfgmod.synth_weights(text)
    x = cfgmod.synth_images(B, 3, hw, hw, seed=5)
    t = cfgmod.synth_truth(B, 90, classes=classes, seed=6)
    runs = []
    for r in range(3):
        net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN | L.NET_TRAIN_REF_CPU)
        cost = N.TrainNetworkDatum(net, x, t)
        convs = [i for i in range(net.n) if net.layer_type(i) == "convolutional"]
        runs.append((cost, {i: N.layer_param(net, i, "weight_updates") for i in convs},
                     {i: N.layer_delta(net, i) for i in convs}))
        net.close()
    for r in (1, 2):
        print("run", r, "cost", runs[r][0], "vs", runs[0][0])
        for i in runs[0][1]:
            a, b = runs[0][1][i], runs[r][1][i]
            da, db = runs[0][2][i], runs[r][2][i]
            print(f"  L{i:3d} wu max-rel {np.abs(a - b).max() / (np.abs(a).max() + 1e-30):.3e}  delta max-rel "
                  f"{np.abs(da - db).max() / (np.abs(da).max() + 1e-30):.3e}")


if __name__ == "__main__":
    main()
