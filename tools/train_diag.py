"""Per-layer error table of one training step vs the reference CPU path (diagnostic; run under gpurun)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from darknet_b200 import cfg as cfgmod, lib as L, network as N
from oracle import pyoracle
from tests import netutil
from tests.netutil import rel_l2

name = sys.argv[1] if len(sys.argv) > 1 else "micro"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
flags = L.NET_DEFAULT | L.NET_TRAIN | L.NET_TRAIN_REF_CPU
kw = dict(w=416, h=416) if name != "micro" else {}
text, cpath, wpath = netutil.materialise(name, batch=B, subdivisions=1, **kw)
ref = pyoracle.RefNet(cpath, wpath, train=True)
import torch
L.check(L.lib().yc_set_device(0), "dev")
net = N.LoadNetwork(cpath, wpath, batch=B, flags=flags)
hw = 64 if name == "micro" else 416
classes = 3 if name == "micro" else 80
x = cfgmod.synth_images(B, 3, hw, hw, seed=5)
truth = cfgmod.synth_truth(B, 90, classes=classes, seed=6)
cr = ref.forward_backward(x, truth)
c = N.TrainNetworkDatum(net, x, truth)
print("cost", c, cr)
for i in range(net.n):
    t = net.layer_type(i)
    row = [f"L{i:3d} {t:14s} out {rel_l2(net.layer_output(i), ref.layer_output(i)):.2e}"]
    d = N.layer_delta(net, i)
    dr = ref.layer_buf(i, "delta")
    if dr is not None:
        row.append(f"delta {rel_l2(d, dr.reshape(d.shape)):.2e} |d| {np.abs(dr).max():.2e}")
    if t == "convolutional":
        row.append(f"wu {rel_l2(N.layer_param(net, i, 'weight_updates'), ref.layer_buf(i, 'weight_updates')):.2e}")
        if net.layer_info(i, 9):
            row.append(f"su {rel_l2(N.layer_param(net, i, 'scale_updates'), ref.layer_buf(i, 'scale_updates')):.2e}")
            row.append(f"var {rel_l2(N.layer_param(net, i, 'variance'), ref.layer_buf(i, 'variance')):.2e}")
        else:
            row.append(f"bu {rel_l2(N.layer_param(net, i, 'bias_updates'), ref.layer_buf(i, 'bias_updates')):.2e}")
    print("  ".join(row))
