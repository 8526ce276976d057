"""Committed golden fixtures (tests/golden/*.npz, produced by tests/golden/make_golden.py from the UNMODIFIED
reference CPU path) against (a) the oracle restatement on the CPU and (b) the CUDA path through the C ABI.
Nothing here reads /root/reference or oracle/_ref."""
import ast
import glob
import os

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod
from tests.netutil import materialise, rel_l2, rel_to_max

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "micro_*.npz")))  # (cv_resize.npz: test_cv_resize_cpu.py)


def _load(path):
    g = np.load(path)
    kw = ast.literal_eval(str(g["kw"]))
    text, cpath, wpath = materialise(str(g["name"]), head_bias=float(g["head_bias"]), head_gain=float(g["head_gain"]), **kw)
    assert text == str(g["cfg"]), "cfg generator drifted from the fixture"
    x = cfgmod.synth_images(1, 3, 64, 64)
    assert np.allclose(x, g["x"].astype(np.float32), atol=1e-3), "image generator drifted from the fixture"
    n = len([k for k in g.files if k.startswith("out")])
    return g, text, cpath, wpath, x, n


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_oracle_vs_golden(built, path):
    from oracle import pynet
    g, text, cpath, wpath, x, n = _load(path)
    weights, _ = cfgmod.read_weights(wpath, text)
    outs = pynet.forward(text, weights, x, fuse_bn=True, infer_avx_maxpool=True)
    assert len(outs) == n
    for i, o in enumerate(outs):
        assert rel_to_max(o, g[f"out{i}"]) < 2e-5, f"layer {i}"


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_cuda_vs_golden(gpu, path):
    from darknet_b200 import lib as L, network as N
    g, text, cpath, wpath, x, n = _load(path)
    net = N.LoadNetwork(cpath, wpath, batch=1, flags=L.NET_DEFAULT | L.NET_REF_CPU_MAXPOOL)
    N.NetworkPredict(net, x)
    for i in range(n):
        if not net.has_output(i):
            continue
        o = net.layer_output(i)
        r = g[f"out{i}"]
        assert rel_l2(o, r) < 1e-2 and rel_to_max(o, r) < 2e-2, f"layer {i} ({net.layer_type(i)})"
    # candidate detections: same count up to threshold-margin flips, boxes close
    thresh = float(g["thresh"])
    d = N.GetNetworkBoxes(net, thresh, 0)
    b5 = d.boxes()
    pre7 = g["pre7"]
    margin = np.sum(np.abs(pre7[:, 4] - thresh) < 0.03)
    assert abs(b5.shape[0] - pre7.shape[0]) <= margin + 1
    if b5.shape[0] == pre7.shape[0]:
        assert np.abs(b5[:, :4] - pre7[:, :4]).max() < 2e-2
    net.close()
