"""Host-side cfg reader + shape walker + darknet `.weights` writer/reader (Python mirror).

Mirrors the reference's INI dialect (`src/parser.cpp:59-100`, `src/option_list.cpp:134-153`) and the
shape rules of the layers on the hot path (conv `src/convolutional_layer.cpp:87-95`, maxpool
`src/maxpool_layer.cpp:62-63` with default pad = size-1 `src/parser.cpp:646`, route with
groups/group_id `src/parser.cpp:819-896`, upsample, shortcut, yolo / Gaussian_yolo).
The `.weights` layout follows `src/parser.cpp:1545-1560,1590-1643` (header: int32 major, minor,
revision + uint64 seen; per conv: biases[n], (BN) scales[n], rolling_mean[n], rolling_variance[n],
weights[n*c/g*k*k] in OIHW fp32).

The C++ runtime (csrc/network.cpp) has its own parser; this module exists so that tools, tests and
bench.py can synthesise weights files without a GPU.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np


def parse_cfg(text: str):
    """Return a list of (section_name, dict) in file order. First must be [net]/[network]."""
    sections = []
    cur = None
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line[0] in "#;":
            continue
        if line[0] == "[":
            cur = (line.strip("[]").strip(), {})
            sections.append(cur)
        else:
            if "=" not in line or cur is None:
                raise ValueError(f"bad cfg line: {raw!r}")
            k, v = line.split("=", 1)
            cur[1][k.strip()] = v.strip()
    if not sections or sections[0][0] not in ("net", "network"):
        raise ValueError("first section must be [net]")
    return sections


@dataclass
class LayerShape:
    kind: str
    index: int
    c: int = 0
    h: int = 0
    w: int = 0
    out_c: int = 0
    out_h: int = 0
    out_w: int = 0
    opts: dict = field(default_factory=dict)
    # conv only
    n: int = 0
    size: int = 0
    stride: int = 1
    pad: int = 0
    bn: int = 0
    groups: int = 1
    activation: str = "linear"
    inputs: tuple = ()

    @property
    def nweights(self):
        return self.n * (self.c // self.groups) * self.size * self.size

    @property
    def bflops(self):  # convolutional_layer.cpp:714
        return 2.0 * self.nweights * self.out_h * self.out_w / 1e9


def walk_shapes(text: str):
    """Return (net_opts, [LayerShape]) following the reference's shape propagation."""
    secs = parse_cfg(text)
    net = secs[0][1]
    h, w, c = int(net.get("height", 0)), int(net.get("width", 0)), int(net.get("channels", 0))
    layers = []
    for idx, (name, o) in enumerate(secs[1:]):
        L = LayerShape(kind=name, index=idx, c=c, h=h, w=w, opts=o)
        if name in ("convolutional", "conv"):
            L.n = int(o.get("filters", 1))
            L.size = int(o.get("size", 1))
            L.stride = int(o.get("stride", 1))
            L.groups = int(o.get("groups", 1))
            pad = int(o.get("pad", 0))
            L.pad = L.size // 2 if pad else int(o.get("padding", 0))
            L.bn = int(o.get("batch_normalize", 0))
            L.activation = o.get("activation", "logistic")
            L.out_c = L.n
            L.out_h = (h + 2 * L.pad - L.size) // L.stride + 1
            L.out_w = (w + 2 * L.pad - L.size) // L.stride + 1
            L.inputs = (idx - 1,)
        elif name in ("maxpool", "max"):
            L.stride = int(o.get("stride", 1))
            L.size = int(o.get("size", L.stride))
            L.pad = int(o.get("padding", L.size - 1))
            L.out_c = c
            L.out_h = (h + L.pad - L.size) // L.stride + 1
            L.out_w = (w + L.pad - L.size) // L.stride + 1
            L.inputs = (idx - 1,)
        elif name == "upsample":
            L.stride = int(o.get("stride", 2))
            L.out_c, L.out_h, L.out_w = c, h * L.stride, w * L.stride
            L.inputs = (idx - 1,)
        elif name == "route":
            ids = [int(x) for x in o["layers"].split(",")]
            ids = [i if i >= 0 else idx + i for i in ids]
            g, gid = int(o.get("groups", 1)), int(o.get("group_id", 0))
            L.groups = g
            L.opts = dict(o, group_id=gid)
            L.inputs = tuple(ids)
            L.out_h, L.out_w = layers[ids[0]].out_h, layers[ids[0]].out_w
            L.out_c = sum(layers[i].out_c // g for i in ids)
        elif name == "shortcut":
            f = int(o["from"])
            f = f if f >= 0 else idx + f
            L.inputs = (idx - 1, f)
            L.activation = o.get("activation", "linear")
            L.out_c, L.out_h, L.out_w = c, h, w
        elif name in ("yolo", "Gaussian_yolo"):
            L.out_c, L.out_h, L.out_w = c, h, w
            L.inputs = (idx - 1,)
        else:
            raise ValueError(f"layer type [{name}] is outside the hot path")
        layers.append(L)
        c, h, w = L.out_c, L.out_h, L.out_w
    return net, layers


def synth_weights(text: str, seed: int = 42):
    """Synthetic, well-conditioned weights: per conv W ~ sqrt(2/(k*k*c/g))*U(-1,1)
    (the reference's own init law, convolutional_layer.cpp:478-493), seed+layer index; biases small
    random for heads (0 elsewhere would hide bias bugs) ; BN gamma ~ U(0.8,1.2), rolling_mean ~
    U(-0.1,0.1), rolling_variance ~ U(0.5,1.5).  Returns {layer_index: dict(name -> fp32 array)}."""
    _, layers = walk_shapes(text)
    out = {}
    for L in layers:
        if L.kind not in ("convolutional", "conv"):
            continue
        rng = np.random.Generator(np.random.Philox(key=seed + L.index))
        fan = L.size * L.size * L.c // L.groups
        scale = np.sqrt(2.0 / fan)
        d = {"weights": (scale * rng.uniform(-1, 1, L.nweights)).astype(np.float32)}
        d["biases"] = rng.uniform(-0.1, 0.1, L.n).astype(np.float32)
        if L.bn:
            d["scales"] = rng.uniform(0.8, 1.2, L.n).astype(np.float32)
            d["rolling_mean"] = rng.uniform(-0.1, 0.1, L.n).astype(np.float32)
            d["rolling_variance"] = rng.uniform(0.5, 1.5, L.n).astype(np.float32)
        out[L.index] = d
    return out


def shift_head_objectness(text, weights, bias, gain=None, box_gain=None):
    """Sparse-detection regime: the objectness / class filters of every head are scaled by `gain` so the
    logits spread like a trained net's, the objectness bias is set to `bias` so only a few cells pass.  box_gain (default:
    gain) scales the box-coordinate filters separately: a trained net's tw/th stay within a few units, while gain 40+
    on them gives boxes of e^10 image widths."""
    _, layers = walk_shapes(text)
    for L in layers:
        if L.kind in ("yolo", "Gaussian_yolo"):
            conv = layers[L.index - 1]
            gaussian = L.kind == "Gaussian_yolo"
            classes = int(L.opts.get("classes", 20))
            E = classes + (9 if gaussian else 5)
            nbox = 8 if gaussian else 4
            if gain is not None:
                w = weights[conv.index]["weights"].reshape(conv.n, -1)
                bg = np.float32(gain if box_gain is None else box_gain)
                for a in range(conv.n // E):
                    w[a * E:a * E + nbox] *= bg
                    w[a * E + nbox:(a + 1) * E] *= np.float32(gain)
                weights[conv.index]["weights"] = w.reshape(-1)
            b = weights[conv.index]["biases"]
            for a in range(conv.n // E):
                if bias is not None:
                    b[a * E + (8 if gaussian else 4)] = bias


def write_weights(path: str, text: str, weights: dict, seen: int = 0):
    _, layers = walk_shapes(text)
    with open(path, "wb") as f:
        f.write(struct.pack("<iii", 0, 2, 5))
        f.write(struct.pack("<Q", seen))
        for L in layers:
            if L.index not in weights:
                continue
            d = weights[L.index]
            f.write(d["biases"].astype("<f4").tobytes())
            if L.bn:
                f.write(d["scales"].astype("<f4").tobytes())
                f.write(d["rolling_mean"].astype("<f4").tobytes())
                f.write(d["rolling_variance"].astype("<f4").tobytes())
            f.write(d["weights"].astype("<f4").tobytes())


def read_weights(path: str, text: str):
    _, layers = walk_shapes(text)
    out = {}
    with open(path, "rb") as f:
        f.read(12)
        seen = struct.unpack("<Q", f.read(8))[0]
        for L in layers:
            if L.kind not in ("convolutional", "conv"):
                continue
            d = {"biases": np.frombuffer(f.read(4 * L.n), "<f4").copy()}
            if L.bn:
                for k in ("scales", "rolling_mean", "rolling_variance"):
                    d[k] = np.frombuffer(f.read(4 * L.n), "<f4").copy()
            d["weights"] = np.frombuffer(f.read(4 * L.nweights), "<f4").copy()
            out[L.index] = d
    return out, seen


def synth_images(batch: int, c: int, h: int, w: int, seed: int = 1234):
    """fp32 NCHW in [0,1), Philox counter-based."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    return rng.random((batch, c, h, w), dtype=np.float32)


def synth_truth(batch: int, max_boxes: int = 90, classes: int = 80, seed: int = 7):
    """[B, max_boxes*5] (x,y,w,h,id), zero-terminated list (data.cpp:255-259)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    t = np.zeros((batch, max_boxes, 5), np.float32)
    for b in range(batch):
        k = int(rng.integers(1, 9))
        t[b, :k, 0:2] = rng.uniform(0.1, 0.9, (k, 2))
        t[b, :k, 2:4] = rng.uniform(0.05, 0.5, (k, 2))
        t[b, :k, 4] = rng.integers(0, classes, k)
    return t.reshape(batch, max_boxes * 5)
