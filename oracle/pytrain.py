"""TEST INFRASTRUCTURE — one training step (forward + backward + optional SGD update) of a whole cfg on the CPU.

Restates the reference's ForwardNetwork(train) / BackwardNetwork / UpdateNetwork (network.cpp:101-208) and the
layers' CPU forward/backward (convolutional_layer.cpp:1128-1400, batchnorm_layer.cpp:92-255, maxpool_layer.cpp:214-323,
upsample_layer.c:76-103, route_layer.c:87-122, shortcut_layer.c:145-188, yolo_layer.cpp:388-777) with numpy/torch CPU
convolutions for the GEMMs and oracle/yc_oracle_loss.c for the YOLO loss.

Two modes:
  emulate_bf16=False : plain fp32 — PINNED against the unmodified reference (tests/test_oracle_train_pin.py).
  emulate_bf16=True  : the same arithmetic with a round-to-bf16 at exactly the points where the B200 path stores bf16
                       (input, GEMM operand weights, pre-batchnorm z, layer outputs except the fp32 heads, deltas).
                       The GPU step is compared against THIS (tests/test_train_emul_gpu.py): agreement to 1e-2 then
                       says the only difference between the GPU path and the reference is those rounding points.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.nn.functional as F

from darknet_b200 import cfg as cfgmod
from oracle.pyoracle import O, oracle_lib

KIND = {"iou": 0, "giou": 1, "mse": 2, "diou": 3, "ciou": 4}


def _bf(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).float().numpy()


def _act_grad(pre, out, act):
    """gradient_array on the output (leaky/relu/logistic/linear) or on the saved pre-activation (mish/swish)."""
    d = np.ones_like(out)
    O_ = oracle_lib()
    src = pre if act in ("mish", "swish") else out
    g = np.ascontiguousarray(d, np.float32)
    O_.yo_gradient(np.ascontiguousarray(src, np.float32).reshape(-1), g.size, {"logistic": 0, "relu": 1, "linear": 4,
                                                                                 "leaky": 8, "swish": 16, "mish": 17}[act],
                   g.reshape(-1))
    return g


def yolo_loss(out, truth, anchors, mask, opts, net_w, net_h):
    lib = oracle_lib()
    lib.yo_yolo_loss.restype = C.c_float
    B, Cc, H, W = out.shape
    classes = int(opts["classes"])
    ip = np.array([len(mask), int(opts.get("num", 1)), classes, int(opts.get("max", 90)), net_w, net_h,
                   KIND.get(opts.get("iou_loss", "mse"), 0), KIND.get(opts.get("iou_thresh_kind", "iou"), 0),
                   int(opts.get("focal_loss", 0))], np.int32)
    fp = np.array([float(opts.get("ignore_thresh", .5)), float(opts.get("truth_thresh", 1)), float(opts.get("iou_thresh", 1)),
                   float(opts.get("iou_normalizer", .75)), float(opts.get("cls_normalizer", 1)),
                   float(opts.get("label_smooth_eps", 0)), float(opts.get("max_delta", 3.402823466e+38))], np.float32)
    delta = np.zeros_like(out, dtype=np.float32)
    f32p = lambda a: a.ctypes.data_as(C.c_void_p)
    o = np.ascontiguousarray(out, np.float32)
    t = np.ascontiguousarray(truth, np.float32)
    a = np.ascontiguousarray(anchors, np.float32)
    m = np.ascontiguousarray(mask, np.int32)
    lib.yo_yolo_loss.argtypes = [C.c_void_p] * 4 + [C.c_int] * 3 + [C.c_void_p] * 3
    cost = lib.yo_yolo_loss(f32p(o), f32p(t), f32p(a), f32p(m), B, H, W, f32p(ip), f32p(fp), f32p(delta))
    return delta, float(cost)


class Step:
    """State of one network across training steps (weights, momentum accumulators)."""

    def _ref_net(self):
        """the compiled reference's network of the same cfg in training mode (only its loss layers are used)"""
        if getattr(self, "_ref", None) is None:
            import os
            import tempfile
            from oracle import pyoracle
            fd, path = tempfile.mkstemp(suffix=".cfg", prefix="yc_walker_")
            with os.fdopen(fd, "w") as f:
                f.write(self.text)
            self._ref = pyoracle.RefNet(path, None, train=True)
        return self._ref

    def __init__(self, text, weights, emulate_bf16=False, dgrad_accumulate=False, bn_bias_grad=False):
        self.text = text
        self.net, self.layers = cfgmod.walk_shapes(text)
        self.emul = emulate_bf16
        self.accum = dgrad_accumulate      # cuDNN beta = 1 (convolutional_kernels.cu:735-738) instead of col2im's overwrite
        self.bn_bias = bn_bias_grad        # backward_bias_gpu in BackwardBatchnormLayerGpu (batchnorm_layer.cpp:330)
        self.P = {}
        for L in self.layers:
            if L.kind in ("convolutional", "conv"):
                d = weights[L.index]
                p = {k: np.array(v, np.float32).copy() for k, v in d.items()}
                p["weights"] = p["weights"].reshape(L.n, L.c, L.size, L.size)
                p["weight_updates"] = np.zeros_like(p["weights"])
                p["bias_updates"] = np.zeros(L.n, np.float32)
                p["scale_updates"] = np.zeros(L.n, np.float32)
                self.P[L.index] = p
        self.net_w, self.net_h = int(self.net["width"]), int(self.net["height"])

    def r(self, a):  # a rounding point of the B200 path
        return _bf(a) if self.emul else np.ascontiguousarray(a, np.float32)

    # ------------------------------------------------------------------------------------------------------------
    def forward_backward(self, x, truth):
        self.forward(x, truth)
        return self.backward()

    def adopt_forward(self, outs, zs, means, variances):
        """Teacher forcing: replace the saved forward state by another implementation's (the GPU's) layer outputs,
        pre-batchnorm tensors z and batch statistics, so that backward() differentiates exactly the function values
        that implementation holds (no leaky-branch flips from forward noise)."""
        Ls = self.layers
        for L in Ls:
            i = L.index
            self.out[i] = np.ascontiguousarray(outs[i], np.float32)
        for L in Ls:
            i = L.index
            if L.kind in ("convolutional", "conv"):
                p, s_ = self.P[i], self.saved[i]
                s_["x"] = self.out[i - 1] if i > 0 else s_["x"]
                if L.bn:
                    mean, var = np.asarray(means[i], np.float32), np.asarray(variances[i], np.float32)
                    xc = np.asarray(zs[i], np.float32) - mean[None, :, None, None]
                    xn = xc / np.sqrt(var + np.float32(.000001))[None, :, None, None]
                    pre = (xn * p["scales"][None, :, None, None] + p["biases"][None, :, None, None]).astype(np.float32)
                    s_.update(xc=xc, xn=xn, pre=pre, mean=mean, var=var)
                else:
                    s_["pre"] = self.out[i]   # only leaky/relu/logistic/linear reach here: gradient on the output
            elif L.kind in ("maxpool", "max"):
                _, idx = O.maxpool(self.out[i - 1], L.size, L.stride, L.pad, want_idx=True)
                self.saved[i] = idx

    def forward(self, x, truth):
        Ls = self.layers
        self._truth = truth
        n = len(Ls)
        out, delta, saved = [None] * n, [None] * n, [None] * n
        costs = []
        x0 = self.r(x)
        cur = x0
        for L in Ls:
            i = L.index
            if L.kind in ("convolutional", "conv"):
                p = self.P[i]
                w = self.r(p["weights"])
                z = F.conv2d(torch.from_numpy(cur), torch.from_numpy(w), stride=L.stride, padding=L.pad).numpy()
                head = i + 1 < n and Ls[i + 1].kind in ("yolo", "Gaussian_yolo")
                if L.bn:
                    m = z.shape[0] * z.shape[2] * z.shape[3]
                    mean = z.mean(axis=(0, 2, 3), dtype=np.float64).astype(np.float32)
                    var = (((z - mean[None, :, None, None]) ** 2).sum(axis=(0, 2, 3), dtype=np.float64) / (m - 1)).astype(np.float32)
                    p["rolling_mean"] = .9 * p["rolling_mean"] + .1 * mean
                    p["rolling_variance"] = .9 * p["rolling_variance"] + .1 * var
                    zs = self.r(z)                                                     # z is stored in bf16
                    xc = zs - mean[None, :, None, None]
                    xn = xc / np.sqrt(var + np.float32(.000001))[None, :, None, None]  # blas.c:214
                    pre = (xn * p["scales"][None, :, None, None] + p["biases"][None, :, None, None]).astype(np.float32)
                    y = O.activate(pre, L.activation)
                    saved[i] = dict(xc=xc, xn=xn, pre=pre, mean=mean, var=var)
                else:
                    pre = (z + p["biases"][None, :, None, None]).astype(np.float32)
                    y = O.activate(pre, L.activation)
                    saved[i] = dict(pre=pre)
                cur = y if head else self.r(y)                                         # head convs keep fp32
                saved[i]["x"] = out[i - 1] if i > 0 else x0
                saved[i]["w"] = w
            elif L.kind in ("maxpool", "max"):
                y, idx = O.maxpool(cur, L.size, L.stride, L.pad, want_idx=True)
                saved[i] = idx
                cur = y
            elif L.kind == "upsample":
                cur = O.upsample(cur, L.stride, 1.0)
            elif L.kind == "route":
                g, gid = L.groups, int(L.opts.get("group_id", 0))
                cur = np.ascontiguousarray(np.concatenate(
                    [out[j][:, (out[j].shape[1] // g) * gid:(out[j].shape[1] // g) * (gid + 1)] for j in L.inputs], axis=1))
            elif L.kind == "shortcut":
                cur = self.r(O.shortcut(out[i - 1], out[L.inputs[1]], L.activation))
            elif L.kind in ("yolo", "Gaussian_yolo"):
                mask = [int(v) for v in L.opts["mask"].split(",")]
                dec = O.yolo_decode(out[i - 1], len(mask), int(L.opts["classes"]), float(L.opts.get("scale_x_y", 1)),
                                    gaussian=L.kind == "Gaussian_yolo")
                cur = dec
            else:
                raise ValueError(L.kind)
            out[i] = cur
        self.out, self.saved = out, saved

    def backward(self):
        Ls = self.layers
        n = len(Ls)
        out, saved, truth = self.out, self.saved, self._truth
        delta = [None] * n
        costs = []
        for L in Ls:  # the loss half of ForwardYoloLayer, on the decoded heads
            if L.kind == "yolo":
                mask = [int(v) for v in L.opts["mask"].split(",")]
                anchors = [float(v) for v in L.opts["anchors"].split(",")]
                delta[L.index], cost = yolo_loss(out[L.index], truth, anchors, mask, L.opts, self.net_w, self.net_h)
                costs.append(cost)
            elif L.kind == "Gaussian_yolo":
                # no restatement of the Gaussian loss exists: the UNMODIFIED reference's own layer runs on the head
                # logits this walker holds (ForwardGaussianYoloLayer(train), gaussian_yolo_layer.cpp:486-857)
                _, delta[L.index], cost = self._ref_net().layer_forward_train(L.index, out[L.index - 1], truth)
                costs.append(cost)
        for L in Ls:  # network.cpp:108-109: every layer's delta starts at zero
            if delta[L.index] is None:
                delta[L.index] = np.zeros_like(out[L.index], dtype=np.float32)
        # ---- BackwardNetwork (network.cpp:162-186) ----
        for L in reversed(Ls):
            i = L.index
            if L.kind in ("yolo", "Gaussian_yolo"):      # BackwardYoloLayer: prev.delta += delta
                delta[i - 1] = self.r(delta[i - 1] + delta[i])
            elif L.kind in ("convolutional", "conv"):
                p, s = self.P[i], saved[i]
                B_, N_, oh, ow = out[i].shape
                dy = delta[i]
                g = dy * _act_grad(s["pre"], out[i], L.activation)
                if L.bn:
                    m = B_ * oh * ow
                    var, mean = s["var"], s["mean"]
                    S1 = g.sum(axis=(0, 2, 3), dtype=np.float64)
                    S2 = (g * s["xc"]).sum(axis=(0, 2, 3), dtype=np.float64)
                    p["scale_updates"] += (S2 / np.sqrt(var + np.float32(.000001))).astype(np.float32)
                    if self.bn_bias:
                        p["bias_updates"] += S1.astype(np.float32)
                    ga = p["scales"].astype(np.float64)
                    md = ga * S1 * (-1. / np.sqrt((var + np.float32(.00001)).astype(np.float64)))
                    vd = ga * S2 * (-.5 * np.power((var + np.float32(.00001)).astype(np.float64), -1.5))
                    c0 = (ga / (np.sqrt(var.astype(np.float64)) + .00001)).astype(np.float32)
                    c1 = (vd * 2. / m).astype(np.float32)
                    c2 = (md / m).astype(np.float32)
                    dz = g * c0[None, :, None, None] + c1[None, :, None, None] * s["xc"] + c2[None, :, None, None]
                else:
                    dz = g
                    p["bias_updates"] += g.sum(axis=(0, 2, 3), dtype=np.float64).astype(np.float32)
                dz = self.r(dz.astype(np.float32))
                delta[i] = dz
                xin = torch.from_numpy(s["x"])
                dzt = torch.from_numpy(dz)
                wt = torch.from_numpy(s["w"])
                gw = torch.nn.grad.conv2d_weight(xin, wt.shape, dzt, stride=L.stride, padding=L.pad).numpy()
                p["weight_updates"] += gw
                if i > 0:
                    gx = torch.nn.grad.conv2d_input(xin.shape, wt, dzt, stride=L.stride, padding=L.pad).numpy()
                    delta[i - 1] = self.r(delta[i - 1] + gx) if self.accum else self.r(gx)   # col2im.c:70 overwrites
            elif L.kind in ("maxpool", "max"):
                idx = saved[i]
                d = delta[i - 1].reshape(-1).copy()
                np.add.at(d, idx.reshape(-1), delta[i].reshape(-1))
                delta[i - 1] = self.r(d.reshape(delta[i - 1].shape))
            elif L.kind == "upsample":
                s_ = L.stride
                B_, C_, H_, W_ = delta[i - 1].shape
                add = delta[i].reshape(B_, C_, H_, s_, W_, s_).sum(axis=(3, 5))
                delta[i - 1] = self.r(delta[i - 1] + add)
            elif L.kind == "route":
                g_, gid = L.groups, int(L.opts.get("group_id", 0))
                off = 0
                for j in L.inputs:
                    part = out[j].shape[1] // g_
                    if delta[j] is not None and Ls[j].kind != "yolo":
                        dj = delta[j].copy()
                        dj[:, part * gid:part * (gid + 1)] += delta[i][:, off:off + part]
                        delta[j] = self.r(dj)
                    off += part
            elif L.kind == "shortcut":
                d = delta[i] * _act_grad(out[i], out[i], L.activation) if L.activation != "linear" else delta[i]
                delta[i] = self.r(d)
                delta[i - 1] = self.r(delta[i - 1] + d)
                f = L.inputs[1]
                delta[f] = self.r(delta[f] + d)
        self.delta = delta
        self.cost = float(np.mean(costs)) if costs else 0.0
        return self.cost

    def update(self, lr=None, momentum=None, decay=None):
        """UpdateNetwork -> UpdateConvolutionalLayer (convolutional_layer.cpp:1382-1400)."""
        lr = float(self.net.get("learning_rate", .001)) if lr is None else lr
        momentum = float(self.net.get("momentum", .9)) if momentum is None else momentum
        decay = float(self.net.get("decay", .0001)) if decay is None else decay
        batch = int(self.net.get("batch", 1)) * int(self.net.get("subdivisions", 1))
        for i, p in self.P.items():
            p["weight_updates"] += np.float32(-decay * batch) * p["weights"]
            p["weights"] += np.float32(lr / batch) * p["weight_updates"]
            p["weight_updates"] *= np.float32(momentum)
            p["biases"] += np.float32(lr / batch) * p["bias_updates"]
            p["bias_updates"] *= np.float32(momentum)
            if "scales" in p:
                p["scales"] += np.float32(lr / batch) * p["scale_updates"]
                p["scale_updates"] *= np.float32(momentum)
