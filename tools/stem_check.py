"""Compare the head outputs of a net with and without YC_NET_STEM_IM2COL (same weights, same images)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N

name = sys.argv[1] if len(sys.argv) > 1 else "yolov4"
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 608
B = int(sys.argv[3]) if len(sys.argv) > 3 else 2
text = cfgs.CONFIGS[name](w=hw, h=hw)
weights = cfgmod.synth_weights(text)
cfgmod.shift_head_objectness(text, weights, -6.0, 40.0)
x = cfgmod.synth_images(B, 3, hw, hw, seed=1234)
outs = {}
for tag, flags in (("stem", L.NET_DEFAULT), ("plain", L.NET_DEFAULT & ~L.NET_STEM_IM2COL)):
    net = N.LoadSynthetic(text, weights, batch=B, flags=flags)
    N.NetworkPredict(net, x)
    outs[tag] = {i: np.array(net.layer_output(i), copy=True) for i in range(net.n) if net.has_output(i) and i in (0, 1, 2, 10, 50, 100) or net.layer_type(i) == "yolo"}
    rows, counts = N.DetectBatch(net, 0.25, 0.45)
    print(tag, "detections", counts.tolist())
    net.close()
for i in outs["stem"]:
    a, b = outs["stem"][i], outs["plain"][i]
    print(f"layer {i}: rel-L2 {np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30):.3e}  max|a| {np.abs(a).max():.3f}")
