"""Debug: run-to-run reproducibility of every layer output (image 0) on the bench workload, per batch size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N
hw = int(sys.argv[1]) if len(sys.argv) > 1 else 608
flags = L.NET_DEFAULT
for a in sys.argv[2:]:
    if a == "plain": flags &= ~L.NET_STEM_IM2COL
    if a == "nograph": flags &= ~L.NET_CUDA_GRAPH
    if a == "nofuse": flags &= ~L.NET_FUSE_SHORTCUT
    if a == "noelide": flags &= ~L.NET_ELIDE_ROUTE
text = cfgs.CONFIGS["yolov4"](w=hw, h=hw)
weights = cfgmod.synth_weights(text)
cfgmod.shift_head_objectness(text, weights, -6.0, 40.0)
x = cfgmod.synth_images(16, 3, hw, hw, seed=1234)
base = {}
for B in (2, 16):
    for rep in range(3):
        net = N.LoadSynthetic(text, weights, batch=B, flags=flags)
        for call in range(2):
            N.NetworkPredict(net, x[:B])
            o = {i: np.array(net.layer_output(i)[0], copy=True) for i in range(net.n) if net.has_output(i)}
            if B not in base:
                base[B] = o
                continue
            bad = [(i, float(np.abs(o[i] - base[B][i]).max())) for i in sorted(o) if not np.array_equal(o[i], base[B][i])]
            print(f"B={B} rep {rep} call {call}: {len(bad)} layers differ from the first run", bad[:3], flush=True)
        net.close()
bad = [(i, float(np.abs(base[2][i] - base[16][i]).max())) for i in sorted(base[2]) if not np.array_equal(base[2][i], base[16][i])]
print("B=2 vs B=16:", len(bad), "layers differ", bad[:4])
