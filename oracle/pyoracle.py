"""TEST INFRASTRUCTURE — ctypes windows onto the two checkers.

* `RefNet`   : the UNMODIFIED reference CPU path (oracle/_ref/libyolo_ref.so, built by oracle/Makefile
               from /root/reference/src + oracle/ref_shim.cpp).  Travels to the GPU box prebuilt.
* `O`        : our own C restatement (oracle/libyc_oracle.so from oracle/yc_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libyolo_ref.so")
ORACLE_SO = os.path.join(HERE, "libyc_oracle.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def build(ref: bool = True):
    """Build the checkers (building the checker is not using it)."""
    subprocess.run(["make", "-C", HERE, "oracle"], check=True, capture_output=True)
    if ref and os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-C", HERE, "-j8", "ref"], check=True, capture_output=True)


def have_ref():
    return os.path.exists(REF_SO)


_ref = None
_orc = None


def ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.ref_load.restype = C.c_void_p
        L.ref_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.ref_load_nofuse.restype = C.c_void_p
        L.ref_load_nofuse.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_num_layers.argtypes = [C.c_void_p]
        L.ref_network_ptr.restype = C.c_void_p
        L.ref_network_ptr.argtypes = [C.c_void_p]
        L.ref_layer_ptr.restype = C.c_void_p
        L.ref_layer_ptr.argtypes = [C.c_void_p, C.c_int]
        L.ref_sizeof.argtypes = [C.c_int]
        L.ref_net_dim.argtypes = [C.c_void_p, C.c_int]
        L.ref_layer_int.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.ref_layer_buf.restype = C.POINTER(C.c_float)
        L.ref_layer_buf.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_longlong)]
        L.ref_predict.restype = C.POINTER(C.c_float)
        L.ref_predict.argtypes = [C.c_void_p, _f32p]
        L.ref_forward.argtypes = [C.c_void_p, _f32p, C.c_void_p, C.c_int]
        L.ref_forward_backward.restype = C.c_float
        L.ref_forward_backward.argtypes = [C.c_void_p, _f32p, _f32p]
        L.ref_train_datum.restype = C.c_float
        L.ref_train_datum.argtypes = [C.c_void_p, _f32p, _f32p]
        L.ref_update.argtypes = [C.c_void_p]
        L.ref_set_hyper.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float]
        L.ref_save_weights.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_get_boxes.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_float, _f32p,
                                    C.c_void_p, C.c_int]
        L.ref_time_predict.restype = C.c_double
        L.ref_time_predict.argtypes = [C.c_void_p, _f32p, C.c_int]
        L.ref_gemm.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, _f32p, C.c_int,
                               _f32p, C.c_int, C.c_float, _f32p, C.c_int]
        L.ref_im2col.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_activate.argtypes = [_f32p, C.c_int, C.c_int, _f32p, _f32p]
        L.ref_activation_id.argtypes = [C.c_char_p]
        L.ref_scale_channels.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, C.c_void_p,
                                         C.c_void_p, C.c_void_p]
        L.ref_reorg.argtypes = [_f32p] + [C.c_int] * 6 + [_f32p]
        L.ref_curr_lr.restype = C.c_float
        L.ref_curr_lr.argtypes = [C.c_void_p, C.c_longlong, C.c_int]
        L.ref_resize.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_layer_forward_train.restype = C.c_float
        L.ref_layer_forward_train.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p, _f32p]
        L.ref_nms.argtypes = [_f32p, _f32p, _f32p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, _i32p]
        _ref = L
    return _ref


class RefNet:
    """The reference network (CPU path). fused=True == LoadNetwork(train=false) (batch forced to 1,
    BN folded, parser.cpp:1114-1115,1866-1867)."""

    def __init__(self, cfg_path: str, weights_path: str | None, train: bool = False, fuse: bool = True):
        L = ref_lib()
        w = (weights_path or "").encode()
        # the reference prints its layer table to stdout/stderr; keep test logs quiet
        devnull = os.open(os.devnull, os.O_WRONLY)
        so, se = os.dup(1), os.dup(2)
        try:
            os.dup2(devnull, 1)
            os.dup2(devnull, 2)
            if fuse and not train:
                self.h = L.ref_load(cfg_path.encode(), w, 0)
            else:
                self.h = L.ref_load_nofuse(cfg_path.encode(), w, 1 if train else 0)
        finally:
            os.dup2(so, 1)
            os.dup2(se, 2)
            os.close(devnull)
            os.close(so)
            os.close(se)
        self.L = L
        self.n = L.ref_num_layers(self.h)
        self.w, self.hh, self.c, self.batch = (L.ref_net_dim(self.h, i) for i in range(4))

    def close(self):
        if self.h:
            self.L.ref_free(self.h)
            self.h = None

    def resize(self, w, h):
        """ResizeNetwork (network.cpp:255-383)."""
        self.L.ref_resize(self.h, w, h)
        self.w, self.hh = (self.L.ref_net_dim(self.h, i) for i in range(2))

    def layer_int(self, i, key):
        return self.L.ref_layer_int(self.h, i, key.encode())

    def layer_buf(self, i, key):
        cnt = C.c_longlong(0)
        p = self.L.ref_layer_buf(self.h, i, key.encode(), C.byref(cnt))
        if not p or cnt.value == 0:
            return None
        return np.ctypeslib.as_array(p, shape=(cnt.value,)).copy()

    def layer_output(self, i):
        """[batch, out_c, out_h, out_w] fp32 copy of layer i's output."""
        o = self.layer_buf(i, "output")
        b = self.layer_int(i, "batch")
        return o.reshape(b, self.layer_int(i, "out_c"), self.layer_int(i, "out_h"), self.layer_int(i, "out_w"))

    def predict(self, x: np.ndarray):
        x = np.ascontiguousarray(x, np.float32).reshape(-1)
        self.L.ref_predict(self.h, x)

    def forward(self, x, truth=None, train=0):
        x = np.ascontiguousarray(x, np.float32).reshape(-1)
        t = None
        if truth is not None:
            self._truth = np.ascontiguousarray(truth, np.float32).reshape(-1)
            t = self._truth.ctypes.data_as(C.c_void_p)
        self.L.ref_forward(self.h, x, t, train)

    def layer_forward_train(self, i, inp, truth):
        """Layer i's own CPU forward (train) on `inp`: returns (output, delta, cost) — the loss half of a YOLO head."""
        shape = (self.layer_int(i, "batch"), self.layer_int(i, "out_c"), self.layer_int(i, "out_h"), self.layer_int(i, "out_w"))
        out, delta = np.zeros(shape, np.float32), np.zeros(shape, np.float32)
        cost = self.L.ref_layer_forward_train(self.h, i, np.ascontiguousarray(inp, np.float32).reshape(-1),
                                              np.ascontiguousarray(truth, np.float32).reshape(-1), out.reshape(-1),
                                              delta.reshape(-1))
        return out, delta, float(cost)

    def forward_backward(self, x, truth):
        x = np.ascontiguousarray(x, np.float32).reshape(-1)
        t = np.ascontiguousarray(truth, np.float32).reshape(-1)
        return float(self.L.ref_forward_backward(self.h, x, t))

    def boxes(self, thresh, nms_thresh=0.45, nms_kind=0, beta=0.6, max_out=200000):
        """GetNetworkBoxes + NmsSort on batch element 0 → (rows[n,7], probs[n,classes])."""
        classes = self.layer_int(self.n - 1, "classes")
        out7 = np.zeros((max_out, 7), np.float32)
        probs = np.zeros((max_out, classes), np.float32)
        n = self.L.ref_get_boxes(self.h, thresh, nms_thresh, nms_kind, beta, out7.reshape(-1),
                                 probs.ctypes.data_as(C.c_void_p), max_out)
        n = min(n, max_out)
        return out7[:n].copy(), probs[:n].copy()

    def time_predict(self, x, iters=1):
        x = np.ascontiguousarray(x, np.float32).reshape(-1)
        return self.L.ref_time_predict(self.h, x, iters)


def ref_nms(boxes, obj, probs, thresh, nms_kind=0, beta=0.6):
    """The reference's NmsSort on arrays: returns (probs after suppression in ORIGINAL order, final order)."""
    boxes = np.ascontiguousarray(boxes, np.float32)
    obj = np.ascontiguousarray(obj, np.float32)
    p = np.ascontiguousarray(probs, np.float32).copy()
    n, classes = p.shape
    order = np.zeros(n, np.int32)
    ref_lib().ref_nms(boxes.reshape(-1), obj, p.reshape(-1), n, classes, thresh, nms_kind, beta, order)
    return p, order


def oracle_lib():
    global _orc
    if _orc is None:
        L = C.CDLL(ORACLE_SO)
        L.yo_activate1.restype = C.c_float
        L.yo_activate1.argtypes = [C.c_float, C.c_int]
        L.yo_activate.argtypes = [_f32p, C.c_longlong, C.c_int]
        L.yo_gradient.argtypes = [_f32p, C.c_longlong, C.c_int, _f32p]
        L.yo_im2col.argtypes = [_f32p] + [C.c_int] * 7 + [_f32p]
        L.yo_col2im.argtypes = [_f32p] + [C.c_int] * 7 + [_f32p]
        L.yo_gemm.argtypes = [C.c_int] * 5 + [C.c_float, _f32p, C.c_int, _f32p, C.c_int, C.c_float, _f32p, C.c_int]
        L.yo_fuse_bn.argtypes = [_f32p, _f32p, _f32p, _f32p, _f32p, C.c_int, C.c_int]
        L.yo_conv_forward.argtypes = [_f32p, _f32p, _f32p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 10 + [_f32p, C.c_void_p]
        L.yo_conv_forward_train.argtypes = [_f32p, _f32p, _f32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 11 + [_f32p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.yo_conv_backward.argtypes = [_f32p, _f32p] + [C.c_void_p] * 7 + [_f32p] + [C.c_int] * 11 + [_f32p, _f32p, C.c_void_p, C.c_void_p]
        L.yo_maxpool.argtypes = [_f32p] + [C.c_int] * 7 + [_f32p, C.c_void_p, C.c_int]
        L.yo_upsample.argtypes = [_f32p] + [C.c_int] * 5 + [C.c_float, _f32p]
        L.yo_shortcut.argtypes = [_f32p, _f32p, C.c_longlong, C.c_int, _f32p]
        L.yo_reorg.argtypes = [_f32p] + [C.c_int] * 6 + [_f32p]
        L.yo_scale_channels.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.yo_scale_channels_forward.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.yo_scale_channels_backward.argtypes = [_f32p, _f32p, C.c_int, _f32p, _f32p, C.c_int, C.c_int, C.c_int, _f32p, _f32p]
        L.yo_yolo_decode.argtypes = [_f32p] + [C.c_int] * 5 + [C.c_float, _f32p]
        L.yo_gaussian_yolo_decode.argtypes = [_f32p] + [C.c_int] * 5 + [C.c_float, _f32p]
        L.yo_yolo_detections.argtypes = [_f32p] + [C.c_int] * 5 + [_f32p, _i32p, C.c_int, C.c_int, C.c_float, _f32p, C.c_int]
        L.yo_nms_sort.argtypes = [_f32p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float]
        L.yo_box_iou.restype = C.c_float
        L.yo_box_iou.argtypes = [_f32p, _f32p]
        _orc = L
    return _orc


ACT = {"logistic": 0, "relu": 1, "linear": 4, "leaky": 8, "swish": 16, "mish": 17}


class O:
    """numpy-friendly wrappers over oracle/yc_oracle.c (fp32 NCHW / OIHW)."""

    @staticmethod
    def conv_forward(x, w, bias, k, stride, pad, act, dil=1, bn=None, want_pre=False):
        L = oracle_lib()
        B, Cc, H, W = x.shape
        N = w.shape[0]
        oh = (H + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        ow = (W + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        out = np.zeros((B, N, oh, ow), np.float32)
        pre = np.zeros_like(out) if want_pre else None
        x = np.ascontiguousarray(x, np.float32)
        w = np.ascontiguousarray(w, np.float32).reshape(-1)
        bias = np.ascontiguousarray(bias, np.float32)
        if bn is not None:
            sc, mu, var = (np.ascontiguousarray(a, np.float32) for a in bn)
            ps = [a.ctypes.data_as(C.c_void_p) for a in (sc, mu, var)]
        else:
            ps = [None, None, None]
        L.yo_conv_forward(x.reshape(-1), w, bias, ps[0], ps[1], ps[2], B, Cc, H, W, N, k, stride, pad, dil,
                          ACT[act] if isinstance(act, str) else act, out.reshape(-1),
                          pre.ctypes.data_as(C.c_void_p) if want_pre else None)
        return (out, pre) if want_pre else out

    @staticmethod
    def conv_forward_train(x, w, biases, k, stride, pad, act, scales=None, rolling=None, dil=1):
        """Training forward. Returns dict(out, z, xnorm, pre, mean, var, rolling_mean, rolling_var)."""
        L = oracle_lib()
        B, Cc, H, W = x.shape
        N = w.shape[0]
        bn = scales is not None
        oh = (H + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        ow = (W + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        out = np.zeros((B, N, oh, ow), np.float32)
        z, xn, pre = np.zeros_like(out), np.zeros_like(out), np.zeros_like(out)
        mean, var = np.zeros(N, np.float32), np.zeros(N, np.float32)
        rm = np.zeros(N, np.float32) if rolling is None else np.ascontiguousarray(rolling[0], np.float32).copy()
        rv = np.zeros(N, np.float32) if rolling is None else np.ascontiguousarray(rolling[1], np.float32).copy()
        sc = np.ascontiguousarray(scales, np.float32) if bn else np.zeros(N, np.float32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        L.yo_conv_forward_train(np.ascontiguousarray(x, np.float32).reshape(-1), np.ascontiguousarray(w, np.float32).reshape(-1),
                                np.ascontiguousarray(biases, np.float32), vp(sc), vp(rm), vp(rv), vp(mean), vp(var),
                                B, Cc, H, W, N, k, stride, pad, dil, ACT[act] if isinstance(act, str) else act, int(bn),
                                out.reshape(-1), vp(z), vp(xn), vp(pre))
        return dict(out=out, z=z, xnorm=xn, pre=pre, mean=mean, var=var, rolling_mean=rm, rolling_var=rv)

    @staticmethod
    def conv_backward(x, w, fwd, delta, k, stride, pad, act, scales=None, dil=1, want_dx=True,
                      weight_updates=None, bias_updates=None, scale_updates=None):
        """BackwardConvolutionalLayer on the buffers of conv_forward_train (`fwd`). Returns dict."""
        L = oracle_lib()
        B, Cc, H, W = x.shape
        N = w.shape[0]
        bn = scales is not None
        d = np.ascontiguousarray(delta, np.float32).copy()
        wu = np.zeros(w.shape, np.float32) if weight_updates is None else np.ascontiguousarray(weight_updates, np.float32).copy()
        bu = np.zeros(N, np.float32) if bias_updates is None else np.ascontiguousarray(bias_updates, np.float32).copy()
        su = np.zeros(N, np.float32) if scale_updates is None else np.ascontiguousarray(scale_updates, np.float32).copy()
        dx = np.zeros(x.shape, np.float32) if want_dx else None
        sc = np.ascontiguousarray(scales, np.float32) if bn else np.zeros(N, np.float32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        L.yo_conv_backward(np.ascontiguousarray(x, np.float32).reshape(-1), np.ascontiguousarray(w, np.float32).reshape(-1),
                           vp(sc), vp(fwd["mean"]), vp(fwd["var"]), vp(fwd["out"]), vp(fwd["pre"]), vp(fwd["z"]), vp(fwd["xnorm"]),
                           d.reshape(-1), B, Cc, H, W, N, k, stride, pad, dil, ACT[act] if isinstance(act, str) else act, int(bn),
                           wu.reshape(-1), bu, vp(su), vp(dx) if want_dx else None)
        return dict(delta=d, weight_updates=wu, bias_updates=bu, scale_updates=su, dx=dx)

    @staticmethod
    def fuse_bn(w, b, scales, mean, var):
        L = oracle_lib()
        w = np.ascontiguousarray(w, np.float32).copy()
        b = np.ascontiguousarray(b, np.float32).copy()
        n = b.shape[0]
        L.yo_fuse_bn(w.reshape(-1), b, np.ascontiguousarray(scales, np.float32), np.ascontiguousarray(mean, np.float32),
                     np.ascontiguousarray(var, np.float32), n, w.size // n)
        return w, b

    @staticmethod
    def activate(x, act):
        y = np.ascontiguousarray(x, np.float32).copy()
        oracle_lib().yo_activate(y.reshape(-1), y.size, ACT[act])
        return y

    @staticmethod
    def maxpool(x, size, stride, pad=None, want_idx=False, infer_avx=False):
        B, Cc, H, W = x.shape
        pad = size - 1 if pad is None else pad
        oh, ow = (H + pad - size) // stride + 1, (W + pad - size) // stride + 1
        out = np.zeros((B, Cc, oh, ow), np.float32)
        idx = np.zeros((B, Cc, oh, ow), np.int32) if want_idx else None
        oracle_lib().yo_maxpool(np.ascontiguousarray(x, np.float32).reshape(-1), B, Cc, H, W, size, stride, pad,
                                out.reshape(-1), idx.ctypes.data_as(C.c_void_p) if want_idx else None, int(infer_avx))
        return (out, idx) if want_idx else out

    @staticmethod
    def upsample(x, stride=2, scale=1.0):
        B, Cc, H, W = x.shape
        out = np.zeros((B, Cc, H * stride, W * stride), np.float32)
        oracle_lib().yo_upsample(np.ascontiguousarray(x, np.float32).reshape(-1), B, Cc, H, W, stride, scale, out.reshape(-1))
        return out

    @staticmethod
    def shortcut(a, b, act="linear"):
        out = np.zeros_like(a, dtype=np.float32)
        oracle_lib().yo_shortcut(np.ascontiguousarray(a, np.float32).reshape(-1), np.ascontiguousarray(b, np.float32).reshape(-1),
                                 a.size, ACT[act], out.reshape(-1))
        return out

    @staticmethod
    def scale_channels(from_out, scale_in, act="linear"):
        """ForwardScaleChannelsLayer: from_out [B,C,H,W], scale_in [B,C] -> act(scale * from_out)."""
        B, Cc, H, W = from_out.shape
        out = np.zeros_like(from_out, dtype=np.float32)
        oracle_lib().yo_scale_channels_forward(np.ascontiguousarray(from_out, np.float32).reshape(-1),
                                               np.ascontiguousarray(scale_in, np.float32).reshape(-1), B, Cc, H * W, ACT[act],
                                               out.reshape(-1))
        return out

    @staticmethod
    def scale_channels_backward(delta, out, from_out, scale_in, act="linear", d_scale_in=None, d_from=None):
        """BackwardScaleChannelsLayer: returns (delta after gradient_array, d_scale_in [B,C], d_from [B,C,H,W]); the two
        gradients ACCUMULATE onto the arrays passed in."""
        B, Cc, H, W = from_out.shape
        d = np.ascontiguousarray(delta, np.float32).copy()
        ds = np.zeros((B, Cc), np.float32) if d_scale_in is None else np.ascontiguousarray(d_scale_in, np.float32).copy()
        df = np.zeros(from_out.shape, np.float32) if d_from is None else np.ascontiguousarray(d_from, np.float32).copy()
        oracle_lib().yo_scale_channels_backward(d.reshape(-1), np.ascontiguousarray(out, np.float32).reshape(-1), ACT[act],
                                                np.ascontiguousarray(from_out, np.float32).reshape(-1),
                                                np.ascontiguousarray(scale_in, np.float32).reshape(-1), B, Cc, H * W,
                                                ds.reshape(-1), df.reshape(-1))
        return d, ds, df

    @staticmethod
    def reorg(x, stride, forward=0):
        """reorg_cpu on [B,C,H,W]: forward=0 (what a [reorg] layer's forward uses) -> [B, C*s*s, H/s, W/s]; forward=1 is the
        inverse scatter."""
        B, Cc, H, W = x.shape
        s2 = stride * stride
        if forward:
            oc, oh, ow = Cc, H, W                       # reorg_cpu is told the SMALL tensor's dims
            out = np.zeros((B, Cc // s2, H * stride, W * stride), np.float32)
        else:
            oc, oh, ow = Cc * s2, H // stride, W // stride
            out = np.zeros((B, oc, oh, ow), np.float32)
        oracle_lib().yo_reorg(np.ascontiguousarray(x, np.float32).reshape(-1), ow, oh, oc, B, stride, forward, out.reshape(-1))
        return out

    @staticmethod
    def yolo_decode(x, n, classes, sxy, gaussian=False):
        B, Cc, H, W = x.shape
        out = np.zeros_like(x, dtype=np.float32)
        f = oracle_lib().yo_gaussian_yolo_decode if gaussian else oracle_lib().yo_yolo_decode
        f(np.ascontiguousarray(x, np.float32).reshape(-1), B, n, classes, H, W, sxy, out.reshape(-1))
        return out

    @staticmethod
    def yolo_detections(pred, b, n, classes, anchors, mask, net_w, net_h, thresh, max_out=200000):
        B, Cc, H, W = pred.shape
        out = np.zeros((max_out, 5 + classes), np.float32)
        cnt = oracle_lib().yo_yolo_detections(np.ascontiguousarray(pred, np.float32).reshape(-1), b, n, classes, H, W,
                                              np.ascontiguousarray(anchors, np.float32), np.ascontiguousarray(mask, np.int32),
                                              net_w, net_h, thresh, out.reshape(-1), max_out)
        return out[:cnt].copy()

    @staticmethod
    def nms_sort(dets, classes, thresh, nms_kind=0, beta=0.6):
        d = np.ascontiguousarray(dets, np.float32).copy()
        oracle_lib().yo_nms_sort(d.reshape(-1), d.shape[0], classes, thresh, nms_kind, beta)
        return d
