"""Per-operator host wrappers over the C ABI, taking torch CUDA tensors in NHWC.

These mirror the reference's per-layer GPU entry points (ForwardConvolutionalLayerGpu,
ForwardMaxpoolLayerGpu, ... — see include/yc_b200.h for file:line) one call at a time, for unit
tests and microbenchmarks.  The whole-network path lives in csrc/network.cu (NetworkPredict).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as L


def _chan_pad(c, dtype):
    e = 4 if dtype == torch.float32 else 8
    return (c + e - 1) // e * e


def to_nhwc(x_nchw: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """fp32 NCHW (device) -> NHWC [n,h,w,ldc] with the pixel pitch padded to 16 bytes."""
    n, c, h, w = x_nchw.shape
    ldc = _chan_pad(c, dtype)
    out = torch.zeros((n, h, w, ldc), dtype=dtype, device=x_nchw.device)
    x = x_nchw.contiguous().float()
    L.check(L.lib().yc_nchw_to_nhwc(x.data_ptr(), L.view(out, n, h, w, c, ldc), L.stream_ptr()), "yc_nchw_to_nhwc")
    return out


def to_nchw(x_nhwc: torch.Tensor, c: int) -> torch.Tensor:
    n, h, w, ldc = x_nhwc.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x_nhwc.device)
    L.check(L.lib().yc_nhwc_to_nchw(L.view(x_nhwc, n, h, w, c, ldc), out.data_ptr(), L.stream_ptr()), "yc_nhwc_to_nchw")
    return out


class Conv:
    """One convolution layer: packed weights + plan over fixed buffers."""

    def __init__(self, x_buf, x_shape, w_oihw, bias, k, stride, pad, act, dil=1, scale=None,
                 out_dtype=torch.bfloat16, residual=None, residual_act="linear", x_ldc=None, x_coff=0,
                 y_buf=None, y_ldc=None, y_coff=0, plan_flags=0, split=None):
        """split=n1: the filter bank is two layers stacked along out_c ([:n1] and [n1:]) that read the same input; the
        kernel writes them to two separate tensors (self.y, self.y2) — yc_conv_plan_create_ex's fused pair."""
        lib = L.lib()
        n, h, w, c = x_shape
        out_c = w_oihw.shape[0]
        assert w_oihw.shape[1] == c and w_oihw.shape[2] == k
        dev = x_buf.device
        self.w32 = w_oihw.contiguous().float().to(dev)
        self.packed = torch.zeros(lib.yc_conv_packed_bytes(out_c, c, k) // 2, dtype=torch.bfloat16, device=dev)
        L.check(lib.yc_conv_pack(self.w32.data_ptr(), out_c, c, k, self.packed.data_ptr(), L.stream_ptr()), "yc_conv_pack")
        self.bias = bias.contiguous().float().to(dev)
        self.scale = None if scale is None else scale.contiguous().float().to(dev)
        eff = dil * (k - 1) + 1
        ho, wo = (h + 2 * pad - eff) // stride + 1, (w + 2 * pad - eff) // stride + 1
        c1 = out_c if split is None else split
        if y_buf is None:
            y_ldc = _chan_pad(c1, out_dtype)
            y_buf = torch.zeros((n, ho, wo, y_ldc), dtype=out_dtype, device=dev)
        self.y = y_buf
        self.out_shape = (n, ho, wo, out_c)
        y2v = None
        if split is not None:
            ld2 = _chan_pad(out_c - split, out_dtype)
            self.y2 = torch.zeros((n, ho, wo, ld2), dtype=out_dtype, device=dev)
            y2v = C.byref(L.view(self.y2, n, ho, wo, out_c - split, ld2))
        x_ldc = x_ldc if x_ldc is not None else x_buf.shape[-1]
        xv = L.view(x_buf, n, h, w, c, x_ldc, x_coff)
        yv = L.view(y_buf, n, ho, wo, c1, y_ldc, y_coff)
        rv = None
        if residual is not None:
            self.residual = residual
            rv = C.byref(L.view(residual, n, ho, wo, out_c, residual.shape[-1]))
        desc = L.ConvDesc(k, stride, pad, dil, L.ACT[act] if isinstance(act, str) else act,
                          L.ACT[residual_act] if isinstance(residual_act, str) else residual_act)
        self.plan = C.c_void_p()
        L.check(lib.yc_conv_plan_create_ex(C.byref(self.plan), C.byref(desc), C.byref(xv), self.packed.data_ptr(),
                                           self.bias.data_ptr(), None if self.scale is None else self.scale.data_ptr(),
                                           rv, C.byref(yv), y2v, plan_flags), "yc_conv_plan_create_ex")

    def info(self):
        lib = L.lib()
        keys = ["block_n", "block_k", "stages", "tiles", "grid", "a_mode", "smem", "variant"]
        return {k: lib.yc_conv_plan_info(self.plan, i) for i, k in enumerate(keys)}

    def __call__(self):
        L.check(L.lib().yc_conv_fprop(self.plan, L.stream_ptr()), "yc_conv_fprop")
        return self.y

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                L.lib().yc_conv_plan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass


class ConvTrain:
    """One convolutional layer in training mode through the operator-level ABI: forward with batch statistics
    (ForwardConvolutionalLayerGpu with state.train) and backward (BackwardConvolutionalLayerGpu)."""

    def __init__(self, x_buf, x_shape, w_oihw, biases, k, stride, pad, act, scales=None, dil=1):
        lib = L.lib()
        n, h, w, c = x_shape
        dev = x_buf.device
        self.dev, self.x_buf, self.x_shape = dev, x_buf, x_shape
        self.k, self.stride, self.pad, self.dil = k, stride, pad, dil
        self.act = L.ACT[act] if isinstance(act, str) else act
        self.N = N = w_oihw.shape[0]
        self.bn = scales is not None
        eff = dil * (k - 1) + 1
        self.ho, self.wo = ho, wo = (h + 2 * pad - eff) // stride + 1, (w + 2 * pad - eff) // stride + 1
        self.w32 = w_oihw.contiguous().float().to(dev)
        self.packed = torch.zeros(lib.yc_conv_packed_bytes(N, c, k) // 2, dtype=torch.bfloat16, device=dev)
        L.check(lib.yc_conv_pack(self.w32.data_ptr(), N, c, k, self.packed.data_ptr(), L.stream_ptr()), "yc_conv_pack")
        self.packed_d = torch.zeros(lib.yc_conv_dgrad_packed_bytes(N, c, k) // 2, dtype=torch.bfloat16, device=dev)
        L.check(lib.yc_conv_pack_dgrad(self.w32.data_ptr(), N, c, k, self.packed_d.data_ptr(), L.stream_ptr()), "yc_conv_pack_dgrad")
        self.biases = biases.contiguous().float().to(dev)
        self.scales = scales.contiguous().float().to(dev) if self.bn else None
        ldy = _chan_pad(N, torch.bfloat16)
        self.z = torch.zeros((n, ho, wo, ldy), dtype=torch.bfloat16, device=dev)   # pre-BN conv output (l->x)
        self.y = torch.zeros((n, ho, wo, ldy), dtype=torch.bfloat16, device=dev)
        self.stats = torch.zeros((2, N), dtype=torch.float32, device=dev)
        self.mean = torch.zeros(N, dtype=torch.float32, device=dev)
        self.var = torch.zeros(N, dtype=torch.float32, device=dev)
        self.rmean = torch.zeros(N, dtype=torch.float32, device=dev)
        self.rvar = torch.zeros(N, dtype=torch.float32, device=dev)
        self.desc = L.ConvDesc(k, stride, pad, dil, L.LINEAR if self.bn else self.act, L.LINEAR)
        xv = L.view(x_buf, n, h, w, c, x_buf.shape[-1])
        self.plan = C.c_void_p()
        out = self.z if self.bn else self.y
        L.check(lib.yc_conv_plan_create(C.byref(self.plan), C.byref(self.desc), C.byref(xv), self.packed.data_ptr(),
                                        None if self.bn else self.biases.data_ptr(), None, None,
                                        C.byref(L.view(out, n, ho, wo, N, ldy))), "yc_conv_plan_create")
        if self.bn:
            L.check(lib.yc_conv_plan_set_stats(self.plan, self.stats[0].data_ptr(), self.stats[1].data_ptr()), "set_stats")
        # backward buffers
        self.cd = lib.yc_conv_dgrad_channels(N)
        self.dz = torch.zeros((n, ho, wo, self.cd), dtype=torch.bfloat16, device=dev)
        self.ws = torch.zeros(lib.yc_bn_bwd_workspace_floats(N), dtype=torch.float32, device=dev)
        self.scale_updates = torch.zeros(N, dtype=torch.float32, device=dev)
        self.bias_updates = torch.zeros(N, dtype=torch.float32, device=dev)
        self.dw_packed = torch.zeros(lib.yc_conv_wgrad_bytes(N, c, k) // 4, dtype=torch.float32, device=dev)
        self.dx = torch.zeros((n, h, w, x_buf.shape[-1]), dtype=torch.bfloat16, device=dev)
        self.wplan = C.c_void_p()
        fdesc = L.ConvDesc(k, stride, pad, dil, L.LINEAR, L.LINEAR)
        L.check(lib.yc_conv_wgrad_plan_create(C.byref(self.wplan), C.byref(fdesc), C.byref(xv),
                                              C.byref(L.view(self.dz, n, ho, wo, self.cd, self.cd)), N,
                                              self.dw_packed.data_ptr()), "yc_conv_wgrad_plan_create")
        self.dz_s1 = self.dz
        hu, wu = ho, wo
        if stride > 1:
            a, b = C.c_int(), C.c_int()
            L.check(lib.yc_conv_dgrad_dilated_size(C.byref(fdesc), h, w, C.byref(a), C.byref(b)), "dilated_size")
            hu, wu = a.value, b.value
            self.dz_s1 = torch.zeros((n, hu, wu, self.cd), dtype=torch.bfloat16, device=dev)
        self.hu, self.wu = hu, wu
        self.dplan = {}
        self.fdesc = fdesc

    def forward(self):
        lib = L.lib()
        n, h, w, c = self.x_shape
        st = L.stream_ptr()
        if self.bn:
            self.stats.zero_()
        L.check(lib.yc_conv_fprop(self.plan, st), "yc_conv_fprop")
        if self.bn:
            ldy = self.z.shape[-1]
            L.check(lib.yc_bn_finalize(self.stats[0].data_ptr(), self.stats[1].data_ptr(), n * self.ho * self.wo,
                                       self.mean.data_ptr(), self.var.data_ptr(), self.rmean.data_ptr(), self.rvar.data_ptr(),
                                       self.N, st), "yc_bn_finalize")
            L.check(lib.yc_bn_act_fwd(L.view(self.z, n, self.ho, self.wo, self.N, ldy), self.mean.data_ptr(), self.var.data_ptr(),
                                      self.scales.data_ptr(), self.biases.data_ptr(), self.act,
                                      L.view(self.y, n, self.ho, self.wo, self.N, ldy), st), "yc_bn_act_fwd")
        return self.y

    def backward(self, dy, accumulate=False, want_dx=True):
        """dy: NHWC bf16 [n,ho,wo,ld] (l->delta).  Returns dx (NHWC bf16) or None."""
        lib = L.lib()
        n, h, w, c = self.x_shape
        st = L.stream_ptr()
        ho, wo, N = self.ho, self.wo, self.N
        dyv = L.view(dy, n, ho, wo, N, dy.shape[-1])
        dzv = L.view(self.dz, n, ho, wo, self.cd, self.cd)
        ldy = self.z.shape[-1]
        if self.bn:
            L.check(lib.yc_bn_bwd(dyv, L.view(self.z, n, ho, wo, N, ldy), self.mean.data_ptr(), self.var.data_ptr(),
                                  self.scales.data_ptr(), self.biases.data_ptr(), self.act, self.scale_updates.data_ptr(),
                                  None, self.ws.data_ptr(), dzv, st), "yc_bn_bwd")
        else:
            L.check(lib.yc_bias_act_bwd(dyv, L.view(self.y, n, ho, wo, N, ldy), self.act, self.bias_updates.data_ptr(),
                                        self.ws.data_ptr(), dzv, st), "yc_bias_act_bwd")
        L.check(lib.yc_conv_wgrad(self.wplan, st), "yc_conv_wgrad")
        if not want_dx:
            return None
        if self.stride > 1:
            L.check(lib.yc_conv_dgrad_dilate(dzv, self.stride, L.view(self.dz_s1, n, self.hu, self.wu, self.cd, self.cd), st),
                    "yc_conv_dgrad_dilate")
        key = bool(accumulate)
        if key not in self.dplan:
            p = C.c_void_p()
            L.check(lib.yc_conv_dgrad_plan_create(C.byref(p), C.byref(self.fdesc),
                                                  C.byref(L.view(self.dz_s1, n, self.hu, self.wu, self.cd, self.cd)),
                                                  self.packed_d.data_ptr(),
                                                  C.byref(L.view(self.dx, n, h, w, c, self.dx.shape[-1])), int(key)),
                    "yc_conv_dgrad_plan_create")
            self.dplan[key] = p
        L.check(lib.yc_conv_fprop(self.dplan[key], st), "yc_conv_dgrad")
        return self.dx

    def weight_updates(self):
        n, h, w, c = self.x_shape
        out = torch.zeros((self.N, c, self.k, self.k), dtype=torch.float32, device=self.dev)
        L.check(L.lib().yc_conv_wgrad_unpack(self.dw_packed.data_ptr(), self.N, c, self.k, out.data_ptr(), L.stream_ptr()),
                "yc_conv_wgrad_unpack")
        return out

    def winfo(self):
        keys = ["block_n", "splits", "stages", "tiles", "grid", "kb_per_split"]
        return {k: L.lib().yc_conv_wgrad_plan_info(self.wplan, i) for i, k in enumerate(keys)}


def maxpool(x, c, size, stride, pad=None, quirk=False):
    n, h, w, ldc = x.shape
    pad = size - 1 if pad is None else pad
    ho, wo = (h + pad - size) // stride + 1, (w + pad - size) // stride + 1
    y = torch.zeros((n, ho, wo, ldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_maxpool_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, ho, wo, c, ldc), size, stride, pad,
                                   int(quirk), L.stream_ptr()), "yc_maxpool_fwd")
    return y


def upsample(x, c, stride=2, scale=1.0):
    n, h, w, ldc = x.shape
    y = torch.zeros((n, h * stride, w * stride, ldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_upsample_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, h * stride, w * stride, c, ldc), stride,
                                    scale, L.stream_ptr()), "yc_upsample_fwd")
    return y


def shortcut(a, b, c, act="linear"):
    n, h, w, ldc = a.shape
    y = torch.zeros_like(a)
    L.check(L.lib().yc_shortcut_fwd(L.view(a, n, h, w, c, ldc), L.view(b, n, h, w, c, b.shape[-1]),
                                    L.view(y, n, h, w, c, ldc), L.ACT[act], L.stream_ptr()), "yc_shortcut_fwd")
    return y


def route_copy(src, c, dst, dst_coff, src_coff=0):
    n, h, w, ldc = src.shape
    L.check(L.lib().yc_route_copy(L.view(src, n, h, w, c, ldc, src_coff), L.view(dst, n, h, w, c, dst.shape[-1], dst_coff),
                                  L.stream_ptr()), "yc_route_copy")
    return dst


def scale_channels(x, s, c, act="linear"):
    n, h, w, ldc = x.shape
    y = torch.zeros_like(x)
    L.check(L.lib().yc_scale_channels_fwd(L.view(x, n, h, w, c, ldc), L.view(s, n, 1, 1, c, s.shape[-1]),
                                          L.view(y, n, h, w, c, ldc), L.ACT[act], L.stream_ptr()), "yc_scale_channels_fwd")
    return y


def reorg(x, c, stride, forward=0):
    n, h, w, ldc = x.shape
    if forward:
        oc, oh, ow = c // (stride * stride), h * stride, w * stride
    else:
        oc, oh, ow = c * stride * stride, h // stride, w // stride
    oldc = _chan_pad(oc, x.dtype)
    y = torch.zeros((n, oh, ow, oldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_reorg_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, oh, ow, oc, oldc), stride, forward,
                                 L.stream_ptr()), "yc_reorg_fwd")
    return y, oc


def scale_channels_bwd(dy, y, x, s, ds, dx, c, act="linear"):
    """BackwardScaleChannelsLayerGpu on NHWC tensors: dy is updated in place by the activation gradient, ds ([n,1,1,ldc]) and
    dx accumulate."""
    n, h, w, ldc = dy.shape
    yv = None if y is None else C.byref(L.view(y, n, h, w, c, y.shape[-1]))
    L.check(L.lib().yc_scale_channels_bwd(L.view(dy, n, h, w, c, ldc), yv, L.ACT[act], L.view(x, n, h, w, c, x.shape[-1]),
                                          L.view(s, n, 1, 1, c, s.shape[-1]), L.view(ds, n, 1, 1, c, ds.shape[-1]),
                                          L.view(dx, n, h, w, c, dx.shape[-1]), L.stream_ptr()), "yc_scale_channels_bwd")


def reorg_bwd(dy, c, stride, reverse=0):
    """BackwardReorgLayerGpu: dy has the forward layer's OUTPUT shape; returns (dx, channels of dx)."""
    n, h, w, ldc = dy.shape
    if reverse:   # forward scattered small -> big, backward gathers big -> small
        oc, oh, ow = c * stride * stride, h // stride, w // stride
    else:
        oc, oh, ow = c // (stride * stride), h * stride, w * stride
    oldc = _chan_pad(oc, dy.dtype)
    dx = torch.zeros((n, oh, ow, oldc), dtype=dy.dtype, device=dy.device)
    L.check(L.lib().yc_reorg_bwd(L.view(dy, n, h, w, c, ldc), L.view(dx, n, oh, ow, oc, oldc), stride, reverse,
                                 L.stream_ptr()), "yc_reorg_bwd")
    return dx, oc


def yolo_decode(x, c, anchors, classes, scale_x_y, gaussian=False):
    n, h, w, ldc = x.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    L.check(L.lib().yc_yolo_decode(L.view(x, n, h, w, c, ldc), out.data_ptr(), anchors, classes, scale_x_y,
                                   int(gaussian), L.stream_ptr()), "yc_yolo_decode")
    return out


def yolo_extract(decoded, anchors_wh, mask, classes, net_w, net_h, thresh, rows_per_image, gaussian=False):
    n, c, h, w = decoded.shape
    dev = decoded.device
    a = torch.as_tensor(anchors_wh, dtype=torch.float32, device=dev).contiguous()
    m = torch.as_tensor(mask, dtype=torch.int32, device=dev).contiguous()
    rows = torch.zeros((n, rows_per_image, 6 + classes), dtype=torch.float32, device=dev)
    counts = torch.zeros((n,), dtype=torch.int32, device=dev)
    L.check(L.lib().yc_yolo_extract(decoded.data_ptr(), n, h, w, len(mask), classes, a.data_ptr(), m.data_ptr(), net_w,
                                    net_h, thresh, int(gaussian), rows.data_ptr(), rows_per_image, counts.data_ptr(),
                                    L.stream_ptr()), "yc_yolo_extract")
    return rows, counts


def yolo_loss(out_nchw, truth, anchors_wh, mask, classes, net_w, net_h, max_boxes=90, total=None, iou_loss="mse",
              iou_thresh_kind="iou", focal_loss=0, ignore_thresh=.5, truth_thresh=1.0, iou_thresh=1.0,
              iou_normalizer=.75, cls_normalizer=1.0, label_smooth_eps=0.0, max_delta=3.402823466e+38, prev_delta=None,
              gaussian=False, uc_normalizer=1.0):
    """ForwardYoloLayer's (or ForwardGaussianYoloLayer's) training half + Backward*YoloLayer on the device.
    Returns (delta_nchw, stats[8])."""
    kinds = {"iou": 0, "giou": 1, "mse": 2, "diou": 3, "ciou": 4}
    n, c, h, w = out_nchw.shape
    dev = out_nchw.device
    a = torch.as_tensor(anchors_wh, dtype=torch.float32, device=dev).contiguous()
    m = torch.as_tensor(mask, dtype=torch.int32, device=dev).contiguous()
    t = torch.as_tensor(truth, dtype=torch.float32, device=dev).contiguous()
    total = a.numel() // 2 if total is None else total
    d = L.YoloLossDesc(len(mask), total, classes, max_boxes, net_w, net_h, kinds[iou_loss], kinds[iou_thresh_kind],
                       focal_loss, ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer,
                       label_smooth_eps, max_delta, 1 if gaussian else 0, uc_normalizer)
    delta = torch.empty_like(out_nchw)
    stats = torch.zeros(L.lib().yc_yolo_loss_stats_floats(n, h, w), dtype=torch.float32, device=dev)
    pv = None
    if prev_delta is not None:
        pv = C.byref(L.view(prev_delta, n, h, w, c, prev_delta.shape[-1]))
    L.check(L.lib().yc_yolo_loss(C.byref(d), out_nchw.data_ptr(), t.data_ptr(), a.data_ptr(), m.data_ptr(), n, h, w,
                                 delta.data_ptr(), pv, stats.data_ptr(), L.stream_ptr()), "yc_yolo_loss")
    return delta, stats[:8]
