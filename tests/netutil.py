"""Shared helpers for the whole-network parity tests (test infrastructure)."""
from __future__ import annotations

import os
import tempfile

import numpy as np

from darknet_b200 import cfg as cfgmod
from darknet_b200 import cfgs
from darknet_b200.cfg import shift_head_objectness  # noqa: F401

_TMP = tempfile.mkdtemp(prefix="yc_nets_")


def materialise(name, seed=42, head_bias=None, head_gain=None, box_gain=None, **kw):
    """Write <name>.cfg and synthetic <name>.weights to a temp dir; returns (cfg_text, cfg_path, weights_path)."""
    text = cfgs.CONFIGS[name](**kw)
    tag = name + "_" + "_".join(f"{k}{v}" for k, v in sorted(kw.items())) + f"_s{seed}_hb{head_bias}_hg{head_gain}_bg{box_gain}"
    cpath = os.path.join(_TMP, tag + ".cfg")
    wpath = os.path.join(_TMP, tag + ".weights")
    if not os.path.exists(wpath):
        with open(cpath, "w") as f:
            f.write(text)
        weights = cfgmod.synth_weights(text, seed=seed)
        if head_bias is not None or head_gain is not None:
            shift_head_objectness(text, weights, head_bias, head_gain, box_gain)
        cfgmod.write_weights(wpath, text, weights)
    return text, cpath, wpath


def rel_to_max(a, b):
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-12))


def rel_l2(a, b):
    return float(np.linalg.norm((a - b).ravel()) / (np.linalg.norm(b.ravel()) + 1e-12))
