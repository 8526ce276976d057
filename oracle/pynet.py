"""TEST INFRASTRUCTURE — whole-network forward built from the C restatement (oracle/yc_oracle.c).

Walks a cfg like the reference's ForwardNetwork (network.cpp:101-114) calling the restated layer
functions (conv = im2col + gemm + bias/BN + activation, maxpool, upsample, route, shortcut, yolo decode),
fp32 NCHW.  Pinned against the UNMODIFIED reference by tests/test_oracle_pin.py.
"""
from __future__ import annotations

import numpy as np

from darknet_b200 import cfg as cfgmod
from oracle.pyoracle import O


def forward(text: str, weights: dict, x: np.ndarray, fuse_bn: bool = True, infer_avx_maxpool: bool = True):
    """Returns the list of per-layer outputs [B, C, H, W] (inference semantics, train=0)."""
    _, layers = cfgmod.walk_shapes(text)
    outs = []
    cur = np.ascontiguousarray(x, np.float32)
    for L in layers:
        if L.kind in ("convolutional", "conv"):
            d = weights[L.index]
            w = d["weights"].reshape(L.n, L.c, L.size, L.size)
            b = d["biases"]
            if L.bn and fuse_bn:  # FuseConvBatchNorm network.cpp:647-682
                w, b = O.fuse_bn(w, b, d["scales"], d["rolling_mean"], d["rolling_variance"])
                w = w.reshape(L.n, L.c, L.size, L.size)
                cur = O.conv_forward(cur, w, b, L.size, L.stride, L.pad, L.activation)
            elif L.bn:
                cur = O.conv_forward(cur, w, b, L.size, L.stride, L.pad, L.activation,
                                     bn=(d["scales"], d["rolling_mean"], d["rolling_variance"]))
            else:
                cur = O.conv_forward(cur, w, b, L.size, L.stride, L.pad, L.activation)
        elif L.kind in ("maxpool", "max"):
            cur = O.maxpool(cur, L.size, L.stride, L.pad, infer_avx=infer_avx_maxpool)
        elif L.kind == "upsample":
            cur = O.upsample(cur, L.stride, 1.0)
        elif L.kind == "route":
            g, gid = L.groups, int(L.opts.get("group_id", 0))
            parts = []
            for i in L.inputs:
                o = outs[i]
                part = o.shape[1] // g
                parts.append(o[:, part * gid:part * (gid + 1)])
            cur = np.ascontiguousarray(np.concatenate(parts, axis=1))
        elif L.kind == "shortcut":
            cur = O.shortcut(outs[L.index - 1], outs[L.inputs[1]], L.activation)
        elif L.kind in ("yolo", "Gaussian_yolo"):
            n = len(L.opts["mask"].split(","))
            cur = O.yolo_decode(outs[L.index - 1], n, int(L.opts["classes"]), float(L.opts.get("scale_x_y", 1)),
                                gaussian=(L.kind == "Gaussian_yolo"))
        else:
            raise ValueError(L.kind)
        outs.append(cur)
    return outs
